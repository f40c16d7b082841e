// oracle/cvshim -- TEST INFRASTRUCTURE: forwards to the single shim header.
#pragma once
#include "core.hpp"
