// Antipodal.h — forwarding header: the reference's src/common/Antipodal.h surface lives in adk_mt.hpp (GPU-backed).
#pragma once
#include "adk_mt.hpp"
