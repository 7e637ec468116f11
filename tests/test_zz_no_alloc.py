"""GPU: lg_filter_matches_static -- the entry point that allocates nothing (key scratch in a static device array of the
library) -- against lg_filter_matches, through the C ABI, on two calls in a row.  Collected last on purpose.
No counterpart in the reference."""
import pytest
import torch

from conftest import pkg

pytestmark = pytest.mark.gpu


def _scores(B, M, N, seed):
    g = torch.Generator().manual_seed(seed)
    s = torch.randn(B, M + 1, N + 1, generator=g)
    s[:, : M // 2, : N // 2] = torch.round(s[:, : M // 2, : N // 2] * 2) / 2  # plenty of exact ties
    return s.cuda().contiguous()


def _outs(B, M, N, fill):
    dev = "cuda"
    return (torch.full((B, M), fill, device=dev, dtype=torch.int64), torch.full((B, N), fill, device=dev, dtype=torch.int64),
            torch.full((B, M), float(fill), device=dev), torch.full((B, N), float(fill), device=dev))


def test_filter_matches_static_equals_workspace_variant_twice():
    C = pkg("_cabi")
    # second call: a different, smaller problem -- the shared array still holds the first call's keys
    cases = [(3, 300, 257, 11, 0.1, None), (2, 130, 97, 12, 0.0, ([100, 130], [97, 5]))]
    for B, M, N, seed, th, lens in cases:
        s = _scores(B, M, N, seed)
        ml = nl = None
        if lens is not None:
            ml = torch.tensor(lens[0], dtype=torch.int32, device="cuda")
            nl = torch.tensor(lens[1], dtype=torch.int32, device="cuda")
        want = C.filter_matches(s, th, ml, nl)
        out = _outs(B, M, N, -7)
        before = torch.cuda.memory_allocated()
        C.filter_matches_static(s, th, out, ml, nl)
        assert torch.cuda.memory_allocated() == before  # the wrapper allocated nothing through torch either
        torch.cuda.synchronize()
        for got, ref, name in zip(out, want, ("m0", "m1", "ms0", "ms1")):
            assert torch.equal(got, ref), f"{name} differs from lg_filter_matches (B={B}, M={M}, N={N})"


def test_filter_matches_static_rejects_sizes_beyond_its_array():
    C = pkg("_cabi")
    B, M, N = 1, C.FILTER_STATIC_MAX_KEYS, 8  # B*(M+N) > LG_FILTER_STATIC_MAX_KEYS
    s = torch.zeros(1, 2, 2, device="cuda")  # never read: the size check comes first
    m0 = torch.empty(B, M, device="cuda", dtype=torch.int64)
    code = C.lib().lg_filter_matches_static(s.data_ptr(), N + 1, None, None, 0.1, m0.data_ptr(), m0.data_ptr(),
                                            s.data_ptr(), s.data_ptr(), B, M, N, None)
    assert code == -4  # LG_ERR_TOO_LARGE
    with pytest.raises(RuntimeError, match="fixed capacity"):
        C._check(code, "lg_filter_matches_static")
