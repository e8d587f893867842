// AntipodalOptimized.h — forwarding header: the reference's src/common/AntipodalOptimized.h surface lives in adk_mt.hpp (GPU-backed).
#pragma once
#include "adk_mt.hpp"
