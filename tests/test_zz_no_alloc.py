"""pb_error_norms_static — the allocation-free variant of pb_error_norms (scratch in a file-scope static device array,
no counterpart in the reference).  Through the C ABI: its output against the existing
entry point's on two calls in a row (the array keeps the previous call's contents), on a second problem and degree sharing
the same array, and the refusal beyond the array's size.  Collected last on purpose."""
import numpy as np
import pytest

import proton_b200 as pb
from proton_b200 import PB

pytestmark = pytest.mark.gpu

# pb_error_norms and the static variant run the same kernels with the same launch and the same summation order, so the
# sums are expected bit-identical; the assertion uses the 1e-8 relative bound of the repo's error-norm tests
TOL = 1e-8


def close(a, b):
    return np.all(np.abs(a - b) <= TOL * np.abs(b) + 1e-300)


def solved_plain(ctx, N, k, fun=PB.FUN_SINSIN):
    ctx.mesh(pb.mesh_init_params(N, N)).assemble(pb.hho_degree_info(k + 1, k), fun=fun)
    assert ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 100000, True))["exit_reason"] == PB.CG_CONVERGED


def test_static_error_norms_match_twice_in_a_row():
    ctx = pb.Context(0)
    solved_plain(ctx, 48, 2)
    ref = ctx.error_norms()
    assert ref[0] > 0 and ref[2] > 0
    first = ctx.error_norms_static()
    second = ctx.error_norms_static()             # the array still holds the first call's contents
    assert close(first, ref) and close(second, ref), (first, second, ref)
    assert np.array_equal(first, second)
    # another degree and right-hand side, then a cut problem, through the same array (stale contents of a LARGER system before)
    solved_plain(ctx, 20, 1, PB.FUN_TANH)
    ref2 = ctx.error_norms()
    assert close(ctx.error_norms_static(), ref2) and close(ctx.error_norms_static(), ref2)
    ctx.mesh(pb.mesh_init_params(24, 24), PB.QUAD_POLY_DUNAVANT)
    ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
    ctx.assemble(pb.hho_degree_info(2, 1), stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
    assert ctx.conjugated_gradient(pb.cg_params(1e-12, 1e8, 100000, True))["exit_reason"] == PB.CG_CONVERGED
    ref3 = ctx.error_norms()
    assert ref3[0] > 0
    assert close(ctx.error_norms_static(), ref3) and close(ctx.error_norms_static(), ref3)
    # the existing entry point is untouched by the calls in between
    assert np.array_equal(ctx.error_norms(), ref3)
    ctx.close()


def test_static_error_norms_two_contexts_share_the_array():
    a, b = pb.Context(0), pb.Context(0)
    solved_plain(a, 32, 2); solved_plain(b, 16, 3)
    ra, rb = a.error_norms(), b.error_norms()
    for _ in range(2):                            # alternating, serialised calls: each one finds the other's leftovers
        assert close(a.error_norms_static(), ra)
        assert close(b.error_norms_static(), rb)
    a.close(); b.close()


def test_static_error_norms_refuses_systems_beyond_the_array():
    assert pb.lib().pb_static_scratch_bytes() == 16 << 20
    ctx = pb.Context(0)
    ctx.mesh(pb.mesh_init_params(512, 512)).assemble(pb.hho_degree_info(3, 2))
    ctx.set_face_solution(np.zeros(ctx.n_rows))
    with pytest.raises(pb.ProtonError) as ei:
        ctx.error_norms_static()
    assert f"[{PB.ERR_UNSUPPORTED}]" in str(ei.value) and "static" in str(ei.value) and "pb_error_norms" in str(ei.value)
    assert ctx.error_norms()[0] > 0               # the allocating entry point still serves it
    # and a state error, not a crash, before any solution exists
    ctx.mesh(pb.mesh_init_params(8, 8)).assemble(pb.hho_degree_info(2, 1))
    with pytest.raises(pb.ProtonError):
        ctx.error_norms_static()
    ctx.close()
