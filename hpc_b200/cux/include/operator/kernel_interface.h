// compatibility path of 0-frameworks/cux/include/operator/kernel_interface.h
#pragma once
#include "cux/kernel_interface.hpp"
