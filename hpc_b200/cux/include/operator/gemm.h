// compatibility path of 0-frameworks/cux/include/operator/gemm.h
#pragma once
#include "cux/ops.hpp"
