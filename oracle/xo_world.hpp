// ORACLE — TEST INFRASTRUCTURE ONLY (see xo_math.hpp header). PARITY UNPINNED.
// Multi-body worlds: filled in below.
#pragma once
#include "xo_physics.hpp"
