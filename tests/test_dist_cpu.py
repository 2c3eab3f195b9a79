"""world_size-2 (and 3) gloo tests of the host side of the multi-GPU path: the strip partition arithmetic that the
halo exchange relies on (pb_strip_layout), exchanged between real processes over torch.distributed/gloo."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import proton_b200 as pb


def _worker(rank, world, port, Nx, Ny, fbs, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rb, re = pb.strip_rows(Ny, world, rank)
        L = pb.strip_layout(Nx, Ny, rb, re, fbs)
        mine = torch.tensor([rb, re] + [L[k] for k in ("cf_begin", "cf_end", "col_lo", "col_hi", "cell_row0", "n_rows", "halo_lo", "halo_hi")], dtype=torch.int64)
        allv = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allv, mine)
        rows = [dict(zip(("rb", "re", "cf_begin", "cf_end", "col_lo", "col_hi", "cell_row0", "n_rows", "halo_lo", "halo_hi"), v.tolist())) for v in allv]
        nfree = (Nx - 1) + (Ny - 1) * (2 * Nx - 1)
        me = rows[rank]
        # the strips tile the cell rows and the free-face range
        assert rows[0]["rb"] == 0 and rows[-1]["re"] == Ny and rows[0]["cf_begin"] == 0 and rows[-1]["cf_end"] == nfree
        if rank + 1 < world:
            nxt = rows[rank + 1]
            assert me["re"] == nxt["rb"] and me["cf_end"] == nxt["cf_begin"]
            # my upper halo is the head of the next rank's owned range, its lower halo is my last face row
            assert me["halo_hi"] == (min(me["cf_end"] + 2 * Nx - 1, nfree) - me["cf_end"]) * fbs
            assert me["halo_hi"] <= nxt["n_rows"]
            assert nxt["col_lo"] >= me["cf_begin"] and nxt["halo_lo"] == (nxt["cf_begin"] - nxt["col_lo"]) * fbs
            assert nxt["cell_row0"] == nxt["rb"] - 1
        if rank == 0:
            assert me["halo_lo"] == 0 and me["cell_row0"] == 0
        if rank == world - 1:
            assert me["halo_hi"] == 0
        # a halo-sized message really travels between neighbours (what ncclSend/Recv does on the box)
        total = torch.tensor([me["n_rows"]], dtype=torch.int64)
        dist.all_reduce(total)
        assert int(total) == nfree * fbs
        if world > 1:
            buf = torch.full((max(me["halo_hi"], 1),), float(rank), dtype=torch.float64)
            if rank + 1 < world:
                dist.send(buf[: rows[rank + 1]["halo_lo"]].contiguous() if rows[rank + 1]["halo_lo"] <= buf.numel() else torch.zeros(rows[rank + 1]["halo_lo"], dtype=torch.float64), dst=rank + 1)
            if rank > 0:
                got = torch.empty(me["halo_lo"], dtype=torch.float64)
                dist.recv(got, src=rank - 1)
                assert got.numel() == me["halo_lo"]
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,Nx,Ny,fbs", [(2, 8, 8, 2), (2, 5, 7, 3), (3, 6, 9, 1)])
def test_strip_layout_across_ranks(world, Nx, Ny, fbs):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + world * 7 + Nx) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, Nx, Ny, fbs, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res


def test_strip_rows_even_split():
    for Ny, world in [(8, 2), (7, 3), (4096, 8), (10, 4)]:
        parts = [pb.strip_rows(Ny, world, r) for r in range(world)]
        assert parts[0][0] == 0 and parts[-1][1] == Ny
        assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
        assert max(e - b for b, e in parts) - min(e - b for b, e in parts) <= 1
