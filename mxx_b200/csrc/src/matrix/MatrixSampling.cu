// MatrixSampling.cu — one translation unit of the matrix layer (see Matrix.cu for the list and MatrixShared.cuh for what crosses units).
#include "MatrixShared.cuh"
#include "sampling.cuh"

#include "MatrixCommon.inl"
#include "MatrixSampling.inl"
