// Forwarder: lets sources written against the reference's include path (<lora_phy/phy.hpp>) build against
// the B200 wrappers unchanged.  The implementation lives in ../lora_phy_b200/phy.hpp.
#pragma once
#include "../lora_phy_b200/phy.hpp"
