// Matrix.cu — single translation unit of the matrix layer (the second of the two entry
// files build.rs compiles: cuda/src/Runtime.cu and cuda/src/matrix/Matrix.cu,
// reference build.rs:57-69).  It includes the parts below, mirroring how the reference's
// Matrix.cu:11-19 includes its MatrixData/Arith/NTT/Decompose/Serde sources; the kernels
// live in csrc/include/{ntt,arith,serde}.cuh.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <exception>
#include <mutex>
#include <new>

#include "gpupoly_internal.cuh"
#include "arith.cuh"
#include "ntt.cuh"
#include "ntt_fast.cuh"
#include "ntt_cluster.cuh"
#include "serde.cuh"
#include "compact.cuh"
#include "sampling.cuh"

#include "MatrixCommon.inl"
#include "MatrixData.inl"
#include "MatrixHostLoad.inl"
#include "MatrixSerde.inl"
#include "MatrixNTT.inl"
#include "MatrixArith.inl"
#include "MatrixDecompose.inl"
#include "MatrixCompact.inl"
#include "MatrixSampling.inl"
