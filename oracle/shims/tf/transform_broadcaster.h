// ORACLE shim: forwards <tf/transform_broadcaster.h> of the reference to oracle/shims/ref_shim.h
#pragma once
#include "ref_shim.h"
