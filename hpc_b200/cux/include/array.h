// compatibility path of 0-frameworks/cux/include/array.h
#pragma once
#include "cux/array4d.hpp"
