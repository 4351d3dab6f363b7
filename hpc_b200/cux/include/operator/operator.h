// compatibility path of 0-frameworks/cux/include/operator/operator.h
#pragma once
#include "cux/operator.hpp"
