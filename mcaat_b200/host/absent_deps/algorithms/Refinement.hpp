#pragma once
// STAND-IN: see core/cft.hpp in this directory.
#include "core/cft.hpp"
