"""NvlsAllReduce without a CUDA device: the communicator stays disabled and declines, exactly like the reference's
TorchSymmMemCommunicator (device_communicators/torch_symm_mem.py:60-131, :165-166) - the caller (GroupCoordinator.
all_reduce, parallel_state.py:648-757) then takes its next method.  No fallback arithmetic lives in this repo."""
import torch

from sglang_b200.collective import NvlsAllReduce


def test_disabled_without_cuda_and_declines():
    comm = NvlsAllReduce(group=None, device="cpu")
    assert comm.disabled and comm.buffer is None and comm.max_size == 0
    x = torch.ones(16, dtype=torch.bfloat16)
    assert comm.should_torch_symm_mem_allreduce(x) is False
    assert comm.all_reduce(x) is None


def test_disabled_on_a_cuda_device_string_when_cuda_is_absent():
    if torch.cuda.is_available():
        return  # covered by tests/test_allreduce_nvls_gpu.py
    comm = NvlsAllReduce(group=None, device="cuda:0")
    assert comm.disabled and comm.all_reduce(torch.ones(8)) is None
