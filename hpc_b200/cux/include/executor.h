// compatibility path of 0-frameworks/cux/include/executor.h
#pragma once
#include "cux/executor.hpp"
