// see PPO_Discrete.h in this directory
#pragma once
#include "../../PPO_MultiDiscrete.h"
