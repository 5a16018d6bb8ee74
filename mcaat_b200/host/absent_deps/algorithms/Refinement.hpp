#pragma once
// TEST INFRASTRUCTURE. See core/cft.hpp in this directory.
#include "core/cft.hpp"
