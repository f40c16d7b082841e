// tests/cpp/mock/KeyFrame.h -- TEST INFRASTRUCTURE: the name the reference includes; see slam_mock.h.
#pragma once
#include "slam_mock.h"
