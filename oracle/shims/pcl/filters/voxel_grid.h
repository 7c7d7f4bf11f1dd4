// ORACLE shim: forwards <pcl/filters/voxel_grid.h> of the reference to oracle/shims/ref_shim.h
#pragma once
#include "ref_shim.h"
