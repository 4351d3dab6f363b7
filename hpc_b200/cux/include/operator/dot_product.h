// compatibility path of 0-frameworks/cux/include/operator/dot_product.h
#pragma once
#include "cux/ops.hpp"
