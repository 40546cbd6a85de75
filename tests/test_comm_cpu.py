"""CPU-side checks of the C ABI's NCCL entry points: the library resolves NCCL at run time, creates a unique id without a GPU,
and rejects bad arguments without touching a device."""
import ctypes as C

import pytest


def test_unique_id_and_contract_errors():
    from llm_fp8_b200 import _lib
    lib = _lib.load()
    a, b = C.create_string_buffer(128), C.create_string_buffer(128)
    assert lib.llmfp8_nccl_unique_id(a) == 0, lib.llmfp8_last_error()
    assert lib.llmfp8_nccl_unique_id(b) == 0
    assert a.raw != b.raw and any(a.raw)
    assert lib.llmfp8_nccl_unique_id(None) < 0
    comm = C.c_void_p()
    assert lib.llmfp8_nccl_init(a.raw, 3, 2, C.byref(comm)) < 0 and b"rank" in lib.llmfp8_last_error()
    assert lib.llmfp8_nccl_allreduce_max_f32(None, None, 4, None) < 0
    assert lib.llmfp8_nccl_allreduce_sum_bf16(None, None, 4, None) < 0
    assert lib.llmfp8_nccl_destroy(None) == 0


def test_cabi_comm_needs_a_way_to_share_the_id():
    import torch

    from llm_fp8_b200 import dp
    with pytest.raises((RuntimeError, ValueError)):
        dp.CAbiComm(1, 2, torch.device("cpu"))
    with pytest.raises(ValueError):
        dp.CAbiComm(0, 1, torch.device("cpu"), unique_id=b"short")
