// Eppstein.h — forwarding header: the reference's src/common/Eppstein.h surface lives in adk_mt.hpp (GPU-backed).
#pragma once
#include "adk_mt.hpp"
