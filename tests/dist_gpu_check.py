"""Multi-GPU check, launched by torchrun (one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dist_gpu_check.py
Every rank assembles its strip, the ranks solve together (NCCL halo exchange + all-reduces inside the library), and
rank 0 compares the gathered solution with a whole-mesh single-GPU solve of the same problem."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import proton_b200 as pb
from proton_b200 import PB
from proton_b200.dist import make_rank_context


def main():
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    def global_close(a, b, tol=1e-8):
        """|a - b| <= tol |b| over ALL ranks: a strip that lies outside the cut domain holds an identically zero solution"""
        t = torch.tensor([float(np.sum((a - b) ** 2)), float(np.sum(b ** 2))], dtype=torch.float64, device=f"cuda:{lr}")
        dist.all_reduce(t)
        return float(t[0]) ** 0.5 <= tol * float(t[1]) ** 0.5 + 1e-13

    ok = True
    for (Nx, Ny, k) in [(16, 16, 1), (24, 19, 2), (12, 32, 0), (10, 9, 1)]:      # the last one gives some of 8 ranks a single cell row
        mip = pb.mesh_init_params(Nx, Ny)
        hdi = pb.hho_degree_info(k + 1, k)
        ctx = make_rank_context(mip, PB.QUAD_TENSOR_GAUSS, rank, world, lr, dist)
        ctx.assemble(hdi)
        res = ctx.conjugated_gradient(pb.cg_params(1e-13, 100.0, 10000, True))
        cells, faces = ctx.recover_strip(ctx.row_begin, ctx.row_end)
        gc = [None] * world; gf = [None] * world
        dist.all_gather_object(gc, cells); dist.all_gather_object(gf, faces)
        if rank == 0:
            ref = pb.Context(lr)
            ref.mesh(mip).assemble(hdi)
            r1 = ref.conjugated_gradient(pb.cg_params(1e-13, 100.0, 10000, True))
            full = ref.recover()
            got = np.concatenate(gc + gf)
            err = np.linalg.norm(got - full) / np.linalg.norm(full)
            print(f"N={Nx}x{Ny} k={k} world={world}: iterations {res['iterations']} (1 GPU: {r1['iterations']}), exit {res['exit_reason']}, rel diff {err:.2e}", flush=True)
            ok = ok and err < 1e-9 and res["exit_reason"] == PB.CG_CONVERGED and abs(res["iterations"] - r1["iterations"]) <= 2
            ref.close()
        # multigrid-preconditioned CG (apply_preconditioner = 3) on strips: level 0 distributed, coarser levels replicated
        rm = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e6, 1000, 3))
        cells_m, faces_m = ctx.recover_strip(ctx.row_begin, ctx.row_end)
        gm = [None] * world; dist.all_gather_object(gm, faces_m)
        okm = rm["exit_reason"] == PB.CG_CONVERGED and global_close(faces_m, faces)
        if not okm:
            print(f"rank {rank}: multigrid solve differs (N={Nx}x{Ny} k={k}): {rm}", flush=True)
        ok = ok and okm
        # block-Jacobi extension (apply_preconditioner = 2) on strips: same solution as the diagonal preconditioner
        rb = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e6, 10000, 2))
        cells_b, faces_b = ctx.recover_strip(ctx.row_begin, ctx.row_end)
        okb = rb["exit_reason"] == PB.CG_CONVERGED and global_close(faces_b, faces)
        if not okb:
            print(f"rank {rank}: block-Jacobi solve differs (N={Nx}x{Ny} k={k}): {rb}", flush=True)
        ok = ok and okb
        ctx.close()
    # fictitious-domain problem (cut cells by K2): every rank preprocesses the whole level set, assembles its strip
    for (N, k) in [(20, 1), (24, 2), (256, 2)]:
        mip = pb.mesh_init_params(N, N)
        hdi = pb.hho_degree_info(k + 1, k)
        ctx = make_rank_context(mip, PB.QUAD_POLY_DUNAVANT, rank, world, lr, dist)
        ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
        ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
        res = ctx.conjugated_gradient(pb.cg_params(1e-14, 1e8, 100000, True))
        cells, faces = ctx.recover_strip(ctx.row_begin, ctx.row_end)
        gc = [None] * world; gf = [None] * world
        dist.all_gather_object(gc, cells); dist.all_gather_object(gf, faces)
        if rank == 0:
            ref = pb.Context(lr)
            ref.mesh(mip, PB.QUAD_POLY_DUNAVANT); ref.cut_levelset(0, 0.35, 0.5, 0.5); ref.cut_preprocess(4)
            ref.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)
            r1 = ref.conjugated_gradient(pb.cg_params(1e-14, 1e8, 100000, True))
            full = ref.recover()
            got = np.concatenate(gc + gf)
            err = np.linalg.norm(got - full) / np.linalg.norm(full)
            print(f"fictdom N={N} k={k} world={world}: iterations {res['iterations']} (1 GPU: {r1['iterations']}), exit {res['exit_reason']}, rel diff {err:.2e}", flush=True)
            ok = ok and err < 1e-8 and res["exit_reason"] == PB.CG_CONVERGED
            os.environ["PB_MG_LINES"] = "0"          # the ranks' smoother (the line smoother is the one-rank default): the same operator on any partition
            rm1 = ref.conjugated_gradient(pb.cg_params(1e-13, 1e8, 5000, 3)); e1 = ref.error_norms()[0]
            del os.environ["PB_MG_LINES"]
            ref.close()
        # the same solve with the multigrid preconditioner on the strips: same solution, same H1 error (summed over ranks), same iterations as one GPU
        rm = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 5000, 3))
        cells_m, faces_m = ctx.recover_strip(ctx.row_begin, ctx.row_end)
        et = torch.tensor([ctx.error_norms()[0]], dtype=torch.float64, device=f"cuda:{lr}"); dist.all_reduce(et)
        okm = rm["exit_reason"] == PB.CG_CONVERGED and global_close(faces_m, faces)
        if rank == 0:
            okm = okm and abs(float(et[0]) - e1) <= 1e-8 * e1 and abs(rm["iterations"] - rm1["iterations"]) <= 2
            print(f"fictdom N={N} k={k} world={world} multigrid: iterations {rm['iterations']} (1 GPU: {rm1['iterations']}), H1^2 {float(et[0]):.12e} (1 GPU: {e1:.12e})", flush=True)
        if not okm:
            print(f"rank {rank}: multigrid fictdom solve differs (N={N} k={k}): {rm}", flush=True)
        ok = ok and okm
        # the line smoother (default on one rank only) forced on the strips: columns of horizontal faces cut at the strip boundaries
        os.environ["PB_MG_LINES"] = "1"
        ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_FICTDOM)          # new values generation: the hierarchy is rebuilt
        rl = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 5000, 3))
        del os.environ["PB_MG_LINES"]
        okl = rl["exit_reason"] == PB.CG_CONVERGED and global_close(ctx.face_solution(), faces) and rl["iterations"] <= rm["iterations"] + 2
        if rank == 0:
            print(f"fictdom N={N} k={k} world={world} multigrid, line smoother on strips: iterations {rl['iterations']} (face-block smoother: {rm['iterations']})", flush=True)
        if not okl:
            print(f"rank {rank}: line-smoother solve on strips differs (N={N} k={k}): {rl}", flush=True)
        ok = ok and okl
        ctx.close()
    # two-material interface problem (doubled unknowns on cut faces): every strip owns the contiguous slots of its faces; the cut
    # faces of the strips below are all-reduced inside pb_assemble; halos are slot ranges; H1 error summed over the ranks
    for (N, k) in [(24, 1), (33, 0), (32, 2), (128, 2)]:
        mip = pb.mesh_init_params(N, N)
        hdi = pb.hho_degree_info(k + 1, k)
        ctx = make_rank_context(mip, PB.QUAD_POLY_DUNAVANT, rank, world, lr, dist)
        ctx.cut_levelset(0, 0.35, 0.5, 0.5); ctx.cut_preprocess(4)
        ctx.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
        L = ctx.interface_strip_layout()
        res = ctx.conjugated_gradient(pb.cg_params(1e-13, 1e8, 200000, True))
        faces = ctx.face_solution()
        et = torch.tensor([ctx.error_norms()[0]], dtype=torch.float64, device=f"cuda:{lr}"); dist.all_reduce(et)
        cells_s, faces_s = ctx.recover_strip(ctx.row_begin, ctx.row_end)          # the strip's part of the reference's packed solution vector
        rows_s = ctx.sample_solution()                                             # the strip's rows of interface_uT.dat
        gf = [None] * world; gl = [None] * world; gcs = [None] * world; grs = [None] * world
        dist.all_gather_object(gf, faces); dist.all_gather_object(gl, (L["slot_begin"], L["slot_end"], L["slots"])); dist.all_gather_object(gcs, cells_s)
        dist.all_gather_object(grs, rows_s)
        if rank == 0:
            ref = pb.Context(lr)
            ref.mesh(mip, PB.QUAD_POLY_DUNAVANT); ref.cut_levelset(0, 0.35, 0.5, 0.5); ref.cut_preprocess(4)
            ref.assemble(hdi, stab_kind=PB.STAB_CUT, problem=PB.PROBLEM_INTERFACE)
            r1 = ref.conjugated_gradient(pb.cg_params(1e-13, 1e8, 200000, True))
            x1 = ref.face_solution(); e1 = ref.error_norms()[0]
            got = np.concatenate(gf)
            contiguous = gl[0][0] == 0 and all(gl[i][1] == gl[i + 1][0] for i in range(world - 1)) and gl[-1][1] == gl[-1][2] == ref.n_rows // hdi.fbs
            err = np.linalg.norm(got - x1) / np.linalg.norm(x1) if got.shape == x1.shape else float("inf")
            full1 = ref.recover(); fullN = np.concatenate(gcs + gf)              # cells of all ranks, then faces of all ranks = pb_recover's layout
            errf = np.linalg.norm(fullN - full1) / np.linalg.norm(full1) if fullN.shape == full1.shape else float("inf")
            rows1 = ref.sample_solution(); rowsN = np.concatenate(grs)           # interface_uT.dat: same points bit for bit, same values to the solve
            same_rows = rowsN.shape == rows1.shape and np.array_equal(rowsN[:, :2], rows1[:, :2]) and np.abs(rowsN[:, 2] - rows1[:, 2]).max() <= 1e-10 * np.abs(rows1[:, 2]).max()
            errf = errf if same_rows else float("inf")
            print(f"interface N={N} k={k} world={world}: unknowns {ref.n_rows}, iterations {res['iterations']} (1 GPU: {r1['iterations']}), exit {res['exit_reason']}, "
                  f"rel diff {err:.2e} (recover_strip parts against pb_recover: {errf:.2e}), H1^2 {float(et[0]):.12e} (1 GPU: {e1:.12e})", flush=True)
            ok = ok and contiguous and err < 1e-8 and errf < 1e-8 and res["exit_reason"] == PB.CG_CONVERGED and abs(res["iterations"] - r1["iterations"]) <= 2 and abs(float(et[0]) - e1) <= 1e-8 * e1
            ref.close()
        ctx.close()
    flag = torch.tensor([1 if ok else 0], device=f"cuda:{lr}")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = bool(int(flag))
    dist.destroy_process_group()
    if rank == 0:
        print("DIST CHECK", "PASSED" if ok else "FAILED", flush=True)
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()
