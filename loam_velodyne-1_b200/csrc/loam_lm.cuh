#pragma once
#include "loam_lo.cuh"
namespace loamgpu { struct LMHandle { }; struct TMHost { }; }
