// MatrixDecompose.cu — one translation unit of the matrix layer (see Matrix.cu for the list and MatrixShared.cuh for what crosses units).
#include "MatrixShared.cuh"
#include "ntt_fast.cuh"

#include "MatrixCommon.inl"
#include "MatrixNttPlan.inl"
#include "MatrixDecompose.inl"
#include "MatrixNoAlloc.inl"
