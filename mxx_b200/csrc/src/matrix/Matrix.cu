// Matrix.cu — the matrix layer.  The reference compiles cuda/src/Runtime.cu and cuda/src/matrix/Matrix.cu
// (build.rs:57-69), the latter #including its MatrixData/Arith/NTT/Decompose/Serde sources (Matrix.cu:11-19).
// Here the layer is split into translation units by kernel family so that they compile in parallel (222 kernel
// instantiations; one unit took five minutes):
//
//   MatrixCore.cu              create/destroy/copy, host <-> device RNS batches, compact wire format, big integers
//   MatrixNTT.cu               block NTT / INTT kernels, global radix passes, gpu_matrix_ntt_all / intt_all
//   MatrixNTTCluster.cu        cluster / DSMEM transforms for rings above one shared-memory block
//   MatrixArith.cu             add / sub / mul / mul_scalar / equal / blocks
//   MatrixDecompose.cu         gadget matrices, gadget decomposition (+ fused forward NTT), mul_decompose
//   MatrixDecomposeCluster.cu  decomposition fused with the cluster transform
//   MatrixSampling.cu          samplers, G-lattice sampler, p1 perturbation
//
// A build.rs for this tree lists these files instead of the single Matrix.cu; this file stays as
// the index and compiles to nothing.
