// ORACLE shim: forwards <pcl_conversions/pcl_conversions.h> of the reference to oracle/shims/ref_shim.h
#pragma once
#include "ref_shim.h"
