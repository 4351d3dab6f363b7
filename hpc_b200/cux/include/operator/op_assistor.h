// compatibility path of 0-frameworks/cux/include/operator/op_assistor.h
#pragma once
#include "cux/op_assistor.hpp"
