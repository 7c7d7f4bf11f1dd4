// ORACLE shim: forwards <sensor_msgs/PointCloud2.h> of the reference to oracle/shims/ref_shim.h
#pragma once
#include "ref_shim.h"
