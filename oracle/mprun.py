#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.

mprun: start N processes of one program as the N ranks of a job of the
stand-in MPI in oracle/mpi_stub/mpi.h (the `mpirun -n N` of this repository).

  python oracle/mprun.py -n 8 [--cwd DIR] [--tag-output] -- prog args...

The launcher creates the job's shared-memory segment (a sparse file under
/dev/shm, sized exactly as mpistub_segment_bytes() does), exports
VPB_MPI_RANK / VPB_MPI_SIZE / VPB_MPI_SHM to every rank, waits for all of
them, and kills the job when one rank fails.  Exit code = the first non-zero
rank exit code, else 0.  Each rank also gets LOCAL_RANK (= rank) so that a
CUDA program can pick its device.
"""
import argparse
import os
import subprocess
import sys
import tempfile
import time

HDR_BYTES, COLL_BYTES, RING_BYTES, RING_HDR = 4096, 65536, 4 << 20, 128


def segment_bytes(n):
    return HDR_BYTES + n * COLL_BYTES + n * n * (RING_HDR + RING_BYTES)


def mprun(n, argv, cwd=None, env=None, timeout=None, capture=True, rank_env=None):
    """Run argv as n ranks.  Returns (exit code, [stdout+stderr text per rank])."""
    shm_dir = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
    fd, path = tempfile.mkstemp(prefix="vpbmpi_", dir=shm_dir)
    os.ftruncate(fd, segment_bytes(n))
    os.close(fd)
    procs, outs = [], []
    try:
        for r in range(n):
            e = dict(os.environ if env is None else env)
            e.update(VPB_MPI_RANK=str(r), VPB_MPI_SIZE=str(n), VPB_MPI_SHM=path, LOCAL_RANK=str(r))
            if rank_env:
                e.update(rank_env(r))
            out = tempfile.TemporaryFile(mode="w+") if capture else None
            outs.append(out)
            procs.append(subprocess.Popen(argv, cwd=cwd, env=e, stdout=out, stderr=subprocess.STDOUT if capture else None))
        t0, rc = time.time(), 0
        live = set(range(n))
        while live:
            for r in list(live):
                c = procs[r].poll()
                if c is not None:
                    live.discard(r)
                    if c != 0 and rc == 0:
                        rc = c
            if rc != 0 or (timeout and time.time() - t0 > timeout):
                if rc == 0:
                    rc = 124
                time.sleep(0.2)
                for r in live:
                    procs[r].kill()
                for r in live:
                    procs[r].wait()
                break
            time.sleep(0.01)
        texts = []
        for out in outs:
            if out is None:
                texts.append("")
            else:
                out.seek(0); texts.append(out.read()); out.close()
        return rc, texts
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
        try:
            os.unlink(path)
        except OSError:
            pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-n", type=int, required=True)
    ap.add_argument("--cwd", default=None)
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("--tag-output", action="store_true", help="print every rank's output, prefixed with its rank (default: rank 0 only)")
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    rc, texts = mprun(a.n, cmd, cwd=a.cwd, timeout=a.timeout)
    for r, t in enumerate(texts):
        if a.tag_output or rc != 0:
            for line in t.splitlines():
                print(f"[{r}] {line}")
        elif r == 0:
            sys.stdout.write(t)
    sys.exit(rc)


if __name__ == "__main__":
    main()
