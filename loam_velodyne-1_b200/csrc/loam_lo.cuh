#pragma once
#include "loam_sr.cuh"
namespace loamgpu { struct LOHandle { }; }
