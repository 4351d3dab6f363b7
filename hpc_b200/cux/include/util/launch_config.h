// compatibility path of 0-frameworks/cux/include/util/launch_config.h
#pragma once
#include "cux/launch_config.hpp"
