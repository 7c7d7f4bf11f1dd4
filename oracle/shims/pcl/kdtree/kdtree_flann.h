// ORACLE shim: forwards <pcl/kdtree/kdtree_flann.h> of the reference to oracle/shims/ref_shim.h
#pragma once
#include "ref_shim.h"
