// the reference's driver names this header without the directory in its commented alternative (driver.cpp:4)
#pragma once
#include "../PPO_MultiDiscrete.h"
