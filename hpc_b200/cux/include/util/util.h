// compatibility path of 0-frameworks/cux/include/util/util.h
#pragma once
#include "cux/core.hpp"
