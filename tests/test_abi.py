"""The C-ABI library loads on a machine without a GPU, exports every symbol
include/gpupoly.h declares, covers the reference's extern "C" block, and fails loudly
(no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gpupoly.h")

# the 38 symbols of src/poly/dcrt/gpu.rs:69-240 + gpu_device_mem_info (src/bench_estimator/gpu.rs:10-14)
REFERENCE_ABI = """
gpu_context_create gpu_context_destroy gpu_context_get_N gpu_event_set_wait gpu_event_set_destroy
gpu_matrix_create gpu_matrix_destroy gpu_matrix_copy gpu_matrix_load_rns_batch gpu_matrix_store_rns_batch
gpu_matrix_store_const_coeff_batch gpu_matrix_store_compact_bytes gpu_matrix_load_compact_bytes
gpu_matrix_add gpu_matrix_add_block gpu_matrix_sub gpu_matrix_mul gpu_matrix_equal gpu_matrix_mul_scalar
gpu_matrix_copy_block gpu_matrix_fill_gadget gpu_matrix_fill_small_gadget
gpu_matrix_fill_small_decomposed_identity_chunk gpu_matrix_decompose_base gpu_matrix_decompose_base_small
gpu_matrix_gauss_samp_gq_arb_base gpu_matrix_create_p1_covariance_cache gpu_matrix_destroy_p1_covariance_cache
gpu_matrix_sample_p1_full_cached gpu_matrix_sample_distribution gpu_matrix_sample_distribution_columns
gpu_matrix_ntt_all gpu_matrix_intt_all gpu_device_synchronize gpu_device_count gpu_last_error
gpu_pinned_alloc gpu_pinned_free gpu_device_mem_info
""".split()


def header_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gpu_[A-Za-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def libpath():
    from mxx_b200 import ffi

    if not os.path.exists(ffi.LIB_PATH):
        import __graft_entry__ as ge

        ge.build()
    return ffi.LIB_PATH


def test_header_covers_reference_abi():
    assert len(REFERENCE_ABI) == 39
    declared = set(header_symbols())
    assert not [s for s in REFERENCE_ABI if s not in declared]


def test_library_exports_every_declared_symbol(libpath):
    out = subprocess.check_output(["nm", "-D", "--defined-only", libpath], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    missing = [s for s in header_symbols() if s not in exported]
    assert not missing, missing
    # nothing but the ABI leaks out (static cudart is hidden)
    assert all(s.startswith("gpu_") for s in exported), sorted(s for s in exported if not s.startswith("gpu_"))[:5]


def test_ctypes_binding_matches_header(libpath):
    from mxx_b200 import ffi

    lib = ffi.lib()
    bound = set(ffi.SIGNATURES) | set(ffi.EXTENSIONS)
    assert bound == set(header_symbols())
    for name in bound:
        assert getattr(lib, name)


def test_no_cpu_fallback_without_device(libpath):
    """without a CUDA device every compute entry point fails loudly; nothing computes on the host"""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from mxx_b200 import GpuDCRTPolyParams, GpuError, standard_moduli

    with pytest.raises(GpuError):
        GpuDCRTPolyParams(16, standard_moduli(16, 1, 20), 4)
    from mxx_b200 import ffi

    n = C.c_int(-1)
    assert ffi.lib().gpu_device_count(C.byref(n)) in (0, 1)


def test_product_path_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under mxx_b200/ may import or load it"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "mxx_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in text and "from oracle" not in text and "import oracle" not in text, f
