// compatibility path of 0-frameworks/cux/include/operator/op_factory.h
#pragma once
#include "cux/operator.hpp"
