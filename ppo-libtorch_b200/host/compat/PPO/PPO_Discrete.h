// Include-path shim: the reference's translation units say  #include "PPO/PPO_Discrete.h"  (driver.cpp:5).  With
// -I<repo>/ppo-libtorch_b200/host/compat the reference's driver.cpp compiles UNMODIFIED against this facade.
#pragma once
#include "../../PPO_Discrete.h"
