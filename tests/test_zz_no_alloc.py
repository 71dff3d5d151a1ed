"""-m gpu: gd_scatter_actions_noalloc, the allocation-free variant of gd_step_host's action staging (the landing
buffer is a file-scope static __device__ array of 16 MiB instead of cudaMalloc'ed memory). No counterpart in the
reference: added for callers that cannot allocate (graph capture, serving
loops). Called through the C ABI; its output must equal what the existing entry point (gd_step_host) leaves in the
action tensor, bit for bit (a copy: the tolerance of the byte / index tests), on two calls in a row - the second
one smaller, so stale contents of the shared array would show. Collected last
(test_zz_*)."""
import ctypes as C

import numpy as np
import pytest

from golden_util import unpack
from helpers import engine_params, param_dict, random_actions

pytestmark = pytest.mark.gpu


def test_scatter_actions_noalloc_matches_step_host(tmp_path):
    import torch
    assert torch.cuda.is_available(), "CUDA device required (no CPU fallback exists)"
    from gpudrive_lab_b200 import _capi
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    lib = _capi.load()
    scenes = unpack(["synth_a", "synth_b", "synth_a"], tmp_path)
    pd = param_dict(dynamicsModel=2, maxNumControlledAgents=8)          # DeltaLocal: the step leaves the actions as written
    sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, engine_params(pd), max_agents=128)
    W, S = 3, 128
    ref_rows = sim.action_tensor().to_torch().view(W * S, 10)           # what the existing entry point writes
    out = torch.zeros(W * S, 10, device="cuda")                         # the caller's output buffer
    out[:, 3:] = 7.0                                                     # columns beyond action_cols must stay untouched
    torch.cuda.synchronize()

    def noalloc(src_ptr, rows, cols):
        """The call under test, with the device's free memory read right before and right after it (nothing else
        in between: torch's own temporaries and the engine's first-step set-up allocate, the entry point must not)."""
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        rc = lib.gd_scatter_actions_noalloc(src_ptr, rows, cols, out.data_ptr(), 0)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0, "the entry point must not allocate device memory"
        return rc

    # ---- call 1: all rows, pinned host source
    a1 = np.ascontiguousarray(random_actions(0, W, S, 2, seed=41)[..., :3])
    h1 = torch.from_numpy(a1).pin_memory()
    sim.step_host(h1)                                                   # existing entry point (gd_step_host)
    _capi.check(noalloc(h1.data_ptr(), W * S, 3))
    assert torch.equal(out[:, :3], ref_rows[:, :3])
    assert torch.equal(out[:, :3].cpu(), torch.from_numpy(a1).view(W * S, 3))
    assert bool((out[:, 3:] == 7.0).all())

    # ---- call 2, right after: fewer rows, pageable host source; rows beyond n_rows keep call 1's values
    a2 = np.ascontiguousarray(random_actions(1, W, S, 2, seed=41)[..., :3])
    sim.step_host(torch.from_numpy(a2))
    n2 = 2 * S
    _capi.check(noalloc(a2.ctypes.data, n2, 3))
    assert torch.equal(out[:n2, :3], ref_rows[:n2, :3])
    assert torch.equal(out[n2:, :3].cpu(), torch.from_numpy(a1).view(W * S, 3)[n2:])
    assert bool((out[:, 3:] == 7.0).all())

    # ---- a device source and another column count
    d3 = torch.arange(5 * 2, dtype=torch.float32, device="cuda").view(5, 2)
    _capi.check(noalloc(d3.data_ptr(), 5, 2))
    assert torch.equal(out[:5, :2], d3) and torch.equal(out[:5, 2], ref_rows[:5, 2])

    # ---- sizes beyond the static array are rejected with a clear error, nothing is launched
    rc = lib.gd_scatter_actions_noalloc(h1.data_ptr(), (4 << 20) // 3 + 1, 3, out.data_ptr(), 0)
    assert rc != 0
    msg = lib.gd_last_error().decode()
    assert "maximum" in msg and "4194304" in msg, msg
    assert lib.gd_scatter_actions_noalloc(h1.data_ptr(), 4, 11, out.data_ptr(), 0) != 0
    assert lib.gd_scatter_actions_noalloc(None, 4, 3, out.data_ptr(), 0) != 0
    torch.cuda.synchronize()
