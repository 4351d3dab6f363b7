// compatibility path of 0-frameworks/cux/include/util/data_filler.h
#pragma once
#include "cux/data_filler.hpp"
