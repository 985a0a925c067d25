"""CUDA-graph replay of the launch-bound sequences (utils/graphs.py): a captured sort and a
captured lab3 step must give, on every replay and for new buffer contents, exactly what the
eager path gives."""
import numpy as np
import pytest
import torch

from cuda_mpi_openmp_b200 import _native as N
from cuda_mpi_openmp_b200 import ops
from cuda_mpi_openmp_b200.models import MahalanobisClassifier, RadixSorter, RobertsFilter
from cuda_mpi_openmp_b200.utils import GraphedStep

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.int32, torch.float32])
def test_sort_graph_replays_on_new_contents(cuda_device, dtype):
    n = 50_000 + 3
    keys = torch.zeros(n, dtype=dtype, device=cuda_device)
    step = RadixSorter(cuda_device).graphed(keys)
    launches = N.launch_count()
    for seed in (1, 2, 3):
        g = torch.Generator(device=cuda_device).manual_seed(seed)
        src = torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=cuda_device, generator=g)
        src = src if dtype == torch.int32 else (src.to(torch.float32) * 1e-3)
        keys.copy_(src)
        out = step()
        assert out.data_ptr() == keys.data_ptr()
        assert torch.equal(keys, torch.sort(src).values)
    assert N.launch_count() == launches          # replays go through the graph, not through the C entry points


def test_lab3_graph_replays_on_new_images(cuda_device):
    rng = np.random.default_rng(41)
    h, w, nc = 120, 160, 5
    classes = [np.stack([rng.integers(0, w, 60), rng.integers(0, h, 60)], axis=1) for _ in range(nc)]
    d = torch.zeros((h, w, 4), dtype=torch.uint8, device=cuda_device)
    d.copy_(torch.from_numpy(rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)))
    step = MahalanobisClassifier(cuda_device).graphed(d, classes)
    for _ in range(3):
        img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
        d.copy_(torch.from_numpy(img))
        step()
        params = MahalanobisClassifier.reference_params(img, classes)
        assert np.array_equal(d.cpu().numpy(), MahalanobisClassifier.reference_classify(img, params))


def test_generic_graphed_step_with_roberts(cuda_device):
    rng = np.random.default_rng(42)
    d = torch.zeros((200, 256, 4), dtype=torch.uint8, device=cuda_device)
    out = torch.empty_like(d)
    step = GraphedStep(lambda: ops.roberts(d, out, variant="tma"), device=cuda_device)
    for _ in range(2):
        img = rng.integers(0, 256, size=(200, 256, 4), dtype=np.uint8)
        d.copy_(torch.from_numpy(img))
        assert step() is out
        assert np.array_equal(out.cpu().numpy(), RobertsFilter.reference(img))
