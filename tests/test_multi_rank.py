"""N>1 paths.  CPU: the 2-rank oracle over gloo against the 1-rank oracle (pins
the decomposition / migration / ghost-exchange host logic).  GPU: the CUDA
library on 2 GPUs over NCCL against the 2-rank oracle (needs 2 devices)."""
import os
import socket
import subprocess
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def launch(mode, nproc=2, timeout=900, extra=(), **env):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(ROOT, "tests", "multi_rank_check.py"), "--mode", mode, *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT, env=dict(os.environ, **env))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "PASS" in r.stdout


def test_two_rank_oracle_matches_one_rank_oracle_gloo():
    launch("oracle")


def test_three_rank_oracle_localised_plasma_gloo():
    """3 ranks along an axis, plasma on rank 0 only: neighbour pairs with and without traffic in the same step."""
    launch("oracle", nproc=3, extra=("--localised",))


def mprun_python(nranks, script, timeout=900, **env):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from mprun import mprun
    e = dict(os.environ, MASTER_PORT=str(free_port()), **env)
    rc, texts = mprun(nranks, [sys.executable, os.path.join(ROOT, "tests", script)], cwd=ROOT, env=e, timeout=timeout)
    assert rc == 0, [t[-2000:] for t in texts]
    assert "PASS" in texts[0]


def need_ref(*names):
    for n in names:
        if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", n)):
            pytest.skip(f"oracle/_ref/{n} not built (needs /root/reference at build time)")


@pytest.mark.ref
@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_n_rank_oracle_is_bit_identical_to_the_reference_on_n_ranks(nranks):
    """The reference ITSELF on N ranks (compiled from /root/reference; its MPI is the multi-process
    stand-in oracle/mpi_stub + oracle/mprun.py) against the N-rank oracle, rank by rank, bit for bit,
    after initialize() and after every step: 1-D splits along x, y, z with periodic / absorbing / metal
    walls, and at 4 / 8 ranks the (2,2,1), (2,1,2), (1,2,2) and (2,2,2) bricks (edges and corners)."""
    need_ref("libvpic_ref_scalar.so")
    mprun_python(nranks, "multi_rank_ref_check.py")


@pytest.mark.ref
def test_reference_pcomm_known_answer_deck_passes_on_eight_ranks_of_the_stand_in_mpi(tmp_path):
    """test/integrated/legacy/pcomm.deck (the reference's own 8-rank known-answer test, legacy/CMakeLists.txt:16)
    run by the unmodified reference on the stand-in MPI: validates the stand-in with the reference's own check."""
    need_ref("deckrun_legacy_pcomm_ref")
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from mprun import mprun
    rc, texts = mprun(8, [os.path.join(ROOT, "oracle", "_ref", "deckrun_legacy_pcomm_ref"), "1", "1"], cwd=str(tmp_path), timeout=300)
    assert rc == 0 and all("pass" in t for t in texts), [t[-800:] for t in texts]


@pytest.mark.gpu
@pytest.mark.parametrize("transport", ["put", "put-unfused", "nccl"])
def test_four_gpus_localised_plasma_match_oracle(transport):
    """4 GPUs in a row with all particles starting on rank 0: rank 1 exchanges injectors with rank 0
    while its other neighbour has nothing to send or receive -- whether a pair exchanges a payload must
    follow from that pair's own counts (a rank-local test desynchronised the put transport's sequence
    numbers with 3 or more ranks along an axis)."""
    import torch
    if torch.cuda.device_count() < 4:
        pytest.skip("needs 4 GPUs")
    launch("gpu", nproc=4, extra=("--localised",), VPIC_B200_TRANSPORT=transport)


@pytest.mark.gpu
@pytest.mark.parametrize("transport", ["put", "put-unfused", "nccl"])
def test_two_gpus_match_two_rank_oracle(transport):
    """put: ghost planes and current folds are PACKED by the plane kernels straight
    into the neighbour's IPC-mapped mailbox over NVLink (pack + transfer in one
    kernel, unpack + wait in one kernel), migrating particles are forwarded by
    k_put / k_get; put-unfused: every message through k_put / k_get; nccl: the
    same exchanges through ncclSend/ncclRecv.  Same bit-exact checks either way."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    launch("gpu", VPIC_B200_TRANSPORT=transport)


@pytest.mark.gpu
def test_two_gpus_full_size_box_conservation_properties():
    """2 x (512x256x256 cells, 2^31 particles): global count, Gauss's law residual and
    energy across the ranks with particles migrating every step (put transport)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    launch("gpufull")


@pytest.mark.gpu
@pytest.mark.parametrize("nproc,extra", [(2, ()), (3, ("--localised",)), (4, ()), (8, ())])
def test_ranks_sharing_one_gpu_match_the_n_rank_oracle(nproc, extra):
    """The N-rank path on a box with ONE GPU: N processes, each with its own device mailbox on the same
    device, bootstrapped by exchanging cudaIpc records over gloo (no NCCL: it refuses two ranks on one
    GPU).  Same kernels, same put transport, same bit-exact criteria as the one-GPU-per-rank tests;
    peer stores land in the same HBM instead of crossing NVLink."""
    launch("gpu", nproc=nproc, extra=extra, timeout=1500, VPIC_B200_TRANSPORT="put", VPB_TEST_BOOTSTRAP="handles",
           VPIC_B200_PUT_TIMEOUT_S="120")


@pytest.mark.gpu
@pytest.mark.parametrize("nproc,small", [(2, "16"), (2, "1"), (4, "16")])
def test_ranks_sharing_one_gpu_resident_lazy_sort_match_the_n_rank_oracle(nproc, small):
    """RESIDENT coherence, sort every second step: sort_p orders indices only, advance_p gathers through them and
    leaves the voxel keys for the next sort, and boundary_p -- which removes the emigrants (back-fill) and appends
    the immigrants -- patches those keys instead of invalidating them.  Particle order, voxels and bits against
    the N-rank oracle after every step.  The sorts after the first are segment sorts (stayers ranked inside their
    home segment, strays in the two small regions around them, ordered by one thread per voxel -- or, with
    VPIC_B200_SEG_SMALL=1, by the warp path that crowded voxels take)."""
    launch("gpu", nproc=nproc, extra=("--resident",), timeout=1500, VPIC_B200_TRANSPORT="put", VPB_TEST_BOOTSTRAP="handles",
           VPIC_B200_PUT_TIMEOUT_S="120", VPIC_B200_SEG_SMALL=small)
