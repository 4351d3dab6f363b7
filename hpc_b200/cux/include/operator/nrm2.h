// compatibility path of 0-frameworks/cux/include/operator/nrm2.h
#pragma once
#include "cux/ops.hpp"
