// tests/cpp/mock/MapPoint.h -- TEST INFRASTRUCTURE: the name the reference includes; see slam_mock.h.
#pragma once
#include "slam_mock.h"
