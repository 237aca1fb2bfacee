"""Generates the committed golden fixtures FROM THE REFERENCE ITSELF.  Runs only
where /root/reference and oracle/_ref exist (make -C oracle ref).

  weibel_init.npz     state right after vpic_simulation::initialize() of the
                      reference's own test deck test/unit/energy_comparison/
                      weibel_driver.cc (16x1x1 cells, 2 x 1600 particles,
                      seed_entropy(1)), dumped by oracle/ref_deckrun_main.cc.
  weibel_energies.txt the energies file that run writes over 700 steps; it is
                      byte-identical to the reference's checked-in
                      test/unit/energy_comparison/energies_gold (asserted here).
  kernels_*.npz       inputs/outputs of single kernels on seeded random states,
                      produced by calling the compiled reference.
"""
import os
import struct
import subprocess
import sys
import tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLD = "/root/reference/test/unit/energy_comparison/energies_gold"


def read_dump(path):
    out = {}
    with open(path, "rb") as f:
        assert f.read(8) == b"VPBDUMP1"
        while True:
            nm = f.read(48)
            if len(nm) < 48:
                break
            (n,) = struct.unpack("<Q", f.read(8))
            out[nm.split(b"\0")[0].decode()] = f.read(n)
    return out


def weibel():
    from vpic_b200 import abi
    exe = os.path.join(ROOT, "oracle", "_ref", "deckrun_weibel")
    with tempfile.TemporaryDirectory() as td:
        subprocess.check_call([exe, os.path.join(td, "weibel"), "0"], cwd=td, stdout=subprocess.DEVNULL)
        energies = open(os.path.join(td, "energies")).read()
        assert energies == open(GOLD).read(), "reference build does not reproduce energies_gold"
        d = read_dump(os.path.join(td, "weibel.init.vpb"))
    grid = np.frombuffer(d["grid"], dtype=np.uint8)
    nsp = int(np.frombuffer(d["species.count"], "i4")[0])
    out = dict(grid=grid, misc=np.frombuffer(d["sim.misc"], "i4"),
               field=np.frombuffer(d["field.f"], abi.field_dtype),
               mc=np.frombuffer(d["field.mc"], "f4"), damp=np.frombuffer(d["field.damp"], "f4"),
               interp=np.frombuffer(d["interp.i"], abi.interp_dtype), nsp=np.int32(nsp))
    for k in range(nsp):   # list order = head first (last defined species first)
        out[f"sp{k}_name"] = np.frombuffer(d[f"species{k}.name"], "S1")
        out[f"sp{k}_qm"] = np.frombuffer(d[f"species{k}.qm"], "f4")
        out[f"sp{k}_ints"] = np.frombuffer(d[f"species{k}.ints"], "i4")
        out[f"sp{k}_p"] = np.frombuffer(d[f"species{k}.p"], abi.particle_dtype)
    np.savez_compressed(os.path.join(HERE, "weibel_init.npz"), **out)
    open(os.path.join(HERE, "weibel_energies.txt"), "w").write(energies)
    print("weibel fixtures written:", nsp, "species,", len(out["sp0_p"]), "particles each")


def kernels():
    """Reference outputs of each kernel on seeded random states (small)."""
    from oracle import harness as H
    from common import make_sim, random_particles, random_fields
    rb = H.ref_backend()
    for tag, kw in (("periodic", dict(kind="periodic", n=(6, 5, 4))),
                    ("metal", dict(kind="metal", n=(5, 4, 6))),
                    ("absorbing", dict(kind="absorbing", n=(5, 6, 4), damp=0.01))):
        s = make_sim(rb, H.RefHost(rb.L), sort_interval=3, **kw)
        out = {}
        for k, sp in enumerate(s.species):
            out[f"p{k}_in"] = random_particles(s, sp, 8, 0.2, 100 + k).copy()
        out["f_in"] = random_fields(s, 7).copy()
        # field-only sequence
        s.be.load_interpolator_array(s.ia, s.fa)
        s.be.advance_b(s.fa, 0.5); s.be.synchronize_jf(s.fa); s.be.advance_e(s.fa, 1.0); s.be.advance_b(s.fa, 0.5)
        s.be.load_interpolator_array(s.ia, s.fa)
        out["f_after_fields"] = s.fields().copy()
        out["i_after_fields"] = s.interpolators().copy()
        out["en_after_fields"] = np.array(s.be.energy_f6(s.fa))
        if tag != "absorbing":
            s.initialize()
            for step in range(6):
                s.advance()
            out["f_after_steps"] = s.fields().copy()
            for k, sp in enumerate(s.species):
                out[f"p{k}_after_steps"] = s.particles(sp).copy()
            out["en_after_steps"] = np.array(s.energies())
        np.savez_compressed(os.path.join(HERE, f"kernels_{tag}.npz"), **out)
        print("kernel fixture", tag, "written")


if __name__ == "__main__":
    weibel()
    kernels()
