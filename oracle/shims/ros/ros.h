// ORACLE shim: forwards <ros/ros.h> of the reference to oracle/shims/ref_shim.h
#pragma once
#include "ref_shim.h"
