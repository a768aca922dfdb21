// oracle/ref_driver_multidiscrete.cpp — TEST INFRASTRUCTURE.
// The reference's driver.cpp only runs PPO_Discrete (driver.cpp:4-13, the MultiDiscrete include is
// commented out).  This three-line main runs the reference's PPO_MultiDiscrete::train() unmodified
// so the masked path has a CPU baseline too.
#include <iostream>
#include "PPO/PPO_MultiDiscrete.h"

int main() {
    try {
        PPO_MultiDiscrete algo;
        algo.train();
    } catch (const std::exception& ex) {
        std::cerr << "Error Occured: " << ex.what() << std::endl;
    }
    return 0;
}
