// driver.cpp — the reference's entry point (driver.cpp:4-22) over the B200 facade: `ppo_b200` trains PPO_Discrete from
// ./PPOConfig.toml; `ppo_b200 multidiscrete` runs PPO_MultiDiscrete the same way.
#include <cstring>
#include <iostream>

#include "PPO_MultiDiscrete.h"

int main(int argc, char** argv) {
    try {
        if (argc > 1 && std::strcmp(argv[1], "multidiscrete") == 0) {
            PPO_MultiDiscrete algo;
            algo.train();
        } else {
            PPO_Discrete algo;
            algo.train();
        }
    } catch (const std::exception& e) {
        std::cout << "Exception: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
