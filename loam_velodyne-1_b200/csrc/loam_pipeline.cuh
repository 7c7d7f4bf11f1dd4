#pragma once
#include "loam_lm.cuh"
namespace loamgpu { struct PipelineHandle { }; }
