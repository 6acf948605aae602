"""CPU checks of the C-ABI library: it loads and exports every symbol include/akm.h declares
(no compute calls without a GPU), and argument errors follow the reference's conventions."""
import ctypes

import pytest

from alink_b200 import native


def test_library_exports_every_declared_symbol():
    L = native.lib()
    names = native.declared_symbols()
    assert "akm_assign_accumulate" in names and "akm_predict" in names and len(names) >= 20
    missing = [s for s in names if not hasattr(L, s)]
    assert not missing, missing
    assert b"akm" in L.akm_version()


def test_bad_arguments_are_reported_not_crashed():
    L = native.lib()
    h = ctypes.c_void_p()
    assert L.akm_create(0, 0, native.EUCLIDEAN, ctypes.byref(h)) == -1          # vector size must be positive
    assert b"vector size" in L.akm_last_error(None)
    assert L.akm_create(0, 4, 7, ctypes.byref(h)) == -1                          # distanceType in {EUCLIDEAN, COSINE}
    assert L.akm_create(0, 4, native.EUCLIDEAN, None) == -1


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(native.AkmError) as e:
        native.Handle(4, native.EUCLIDEAN)
    assert e.value.code == -2                                                    # AKM_ERR_CUDA: fails loudly
