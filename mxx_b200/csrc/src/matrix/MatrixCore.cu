// MatrixCore.cu — one translation unit of the matrix layer (see Matrix.cu for the list and MatrixShared.cuh for what crosses units).
#include "MatrixShared.cuh"
#include "serde.cuh"
#include "compact.cuh"

#include "MatrixCommon.inl"
#include "MatrixData.inl"
#include "MatrixHostLoad.inl"
#include "MatrixSerde.inl"
#include "MatrixCompact.inl"
