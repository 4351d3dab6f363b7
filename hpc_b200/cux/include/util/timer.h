// compatibility path of 0-frameworks/cux/include/util/timer.h
#pragma once
#include "cux/timer.hpp"
