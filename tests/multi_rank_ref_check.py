"""Pin the N-rank ORACLE to the N-rank REFERENCE.  Run under oracle/mprun.py (N processes):

  python oracle/mprun.py -n 4 -- python tests/multi_rank_ref_check.py

Every process is rank r of two N-rank programs at once:
  * the compiled reference (oracle/_ref/libvpic_ref_scalar.so) whose MPI is the stand-in of
    oracle/mpi_stub (shared-memory rings between the N processes), i.e. the reference's own
    boundary_p / synchronize_* / ghost exchanges talking to real neighbour ranks, and
  * the C restatement (oracle/liboracle.so) whose exchanges go through gloo,
both driven by the same Python restatement of advance().  After initialize() and after every step
the two must agree BIT FOR BIT on every rank: particle arrays (order included), np, nm, every field
component, interpolators, energies.  This is what makes the N-rank oracle that the GPU tests use
(tests/multi_rank_check.py) a pinned one -- including the multi-axis (2,2,1) / (2,2,2) topologies."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world = int(os.environ["VPB_MPI_RANK"]), int(os.environ["VPB_MPI_SIZE"])
    os.environ.setdefault("RANK", str(rank)); os.environ.setdefault("WORLD_SIZE", str(world))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    import torch.distributed as dist
    dist.init_process_group("gloo")
    from oracle import harness as H
    import multi_rank_check as M
    from common import same_bits
    steps = int(os.environ.get("VPB_CHECK_STEPS", "5"))
    ref = H.ref_backend(simd=False, tpp=1)            # boots the reference: MPI_Init of the stand-in joins the N-rank job
    facts = H.ref_facts(ref.L)
    ok = True
    for kind, gn, topo in M.case_list(world):
        M.LOCALISED_LN = tuple(gn[a] // topo[a] for a in range(3))
        H.set_world(rank, world)
        so, origin, ln = M.build(H.oracle_backend(), kind, gn, topo, rank, 6, 40)
        H.install_gloo_exchange(so)
        sr, _, _ = M.build(ref, kind, gn, topo, rank, 6, 40, host=H.RefHost(ref.L))
        gs, gr = so.grid_struct(so.g), sr.grid_struct(sr.g)
        if list(gs.bc) != list(gr.bc):
            print(f"rank {rank} {kind} {topo}: neighbour ranks differ", list(gs.bc), list(gr.bc)); ok = False
        for s in (so, sr):
            M.with_cleaning(s)
        for step in range(-1, steps):
            if step >= 0:
                so.advance(); sr.advance()
            bad = []
            for k in so.fields().dtype.names:
                if not k.startswith("_") and not same_bits(so.fields()[k], sr.fields()[k]):
                    bad.append(k)
            if not same_bits(so.interpolators(), sr.interpolators()):
                bad.append("interpolators")
            for a, b in zip(so.species, sr.species):
                sa, sb = so.sp_struct(a), sr.sp_struct(b)
                if (sa.np, sa.nm) != (sb.np, sb.nm):
                    bad.append(f"np/nm {sa.np},{sa.nm} vs {sb.np},{sb.nm}")
                elif not same_bits(so.particles(a), sr.particles(b)):
                    bad.append("particles")
            if bad:
                print(f"rank {rank} {kind} {topo} step {step}: oracle != reference: {bad}"); ok = False
                break
        # the reference all-reduces its energies itself (the stand-in adds the ranks in rank order);
        # the oracle returns local values: add them the same way
        mine = so.energies()
        every = [None] * world
        dist.all_gather_object(every, mine)
        eo = [0.0] * len(mine)
        for r in range(world):
            eo = [a + b for a, b in zip(eo, every[r])]
        er = sr.energies()
        if eo != er:
            print(f"rank {rank} {kind} {topo}: energies differ", eo, er); ok = False
        if so.init_report != sr.init_report:
            print(f"rank {rank} {kind} {topo}: initialize() reports differ", so.init_report, sr.init_report); ok = False
        dist.barrier()
    import torch
    flag = torch.tensor([1 if ok else 0]); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("multi_rank_ref_check world", world, "PASS" if flag.item() else "FAIL")
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
