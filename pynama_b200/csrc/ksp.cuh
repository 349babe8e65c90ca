// ksp.cuh
#pragma once
#include "spmv.cuh"
