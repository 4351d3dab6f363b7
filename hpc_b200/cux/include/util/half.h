// compatibility path of 0-frameworks/cux/include/util/half.h
#pragma once
#include "cux/half.hpp"
