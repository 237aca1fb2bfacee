"""Host-side time step: a Python mirror of vpic_simulation::advance()
(reference: src/vpic/advance.cc:15-208) and of the parts of
vpic_simulation::initialize() (src/vpic/initialize.cc:4-64) the hot path needs.

It only sequences calls; every array lives in the POD structs of abi.py and
every computation happens behind a `Backend` (lib.py).  With the product
backend ("b200") all of it runs on the GPU.
"""
import ctypes as C
import numpy as np
from . import abi
from .lib import load_b200, b200_backend

_P = C.c_void_p


class B200Host:
    """Struct constructors from libvpic_b200.so part 3 (pure host code)."""

    def __init__(self):
        self.L = load_b200()

    def new_grid(self):
        return self.L.vpic_b200_new_grid()

    def partition_periodic_box(self, g, lo, hi, n, topo):
        self.L.vpic_b200_partition_periodic_box(g, *map(float, lo), *map(float, hi), *map(int, n), *map(int, topo))

    def partition_metal_box(self, g, lo, hi, n, topo):
        self.L.vpic_b200_partition_metal_box(g, *map(float, lo), *map(float, hi), *map(int, n), *map(int, topo))

    def partition_absorbing_box(self, g, lo, hi, n, topo, pbc):
        self.L.vpic_b200_partition_absorbing_box(g, *map(float, lo), *map(float, hi), *map(int, n), *map(int, topo), int(pbc))

    def set_fbc(self, g, boundary, fbc):
        self.L.vpic_b200_set_fbc(g, boundary, fbc)

    def set_pbc(self, g, boundary, pbc):
        self.L.vpic_b200_set_pbc(g, boundary, pbc)

    def material_list(self, eps=1.0, mu=1.0, sigma=0.0, more=()):
        """"vacuum" (id 0) plus the (name, eps, mu, sigma) materials of `more` (ids 1..)."""
        head = _P(None)
        for name, e, m, sg in (("vacuum", eps, mu, sigma),) + tuple(more):
            self.L.vpic_b200_append_material(self.L.vpic_b200_new_material(name.encode(), e, m, sg), C.byref(head))
        return head.value

    def field_array(self, g, m_list, damp):
        return self.L.new_standard_field_array(g, m_list, damp)

    def interpolator_array(self, g):
        return self.L.vpic_b200_new_interpolator_array(g)

    def accumulator_array(self, g, n_pipeline):
        return self.L.vpic_b200_new_accumulator_array(g, n_pipeline)

    def hydro_array(self, g):
        return self.L.vpic_b200_new_hydro_array(g)

    def species(self, name, q, m, max_np, max_nm, sort_interval, sort_out_of_place, g, host_arrays=True):
        return self.L.vpic_b200_new_species(name.encode(), q, m, max_np, max_nm, sort_interval,
                                            sort_out_of_place, g, 1 if host_arrays else 0)


class Simulation:
    """grid + fields + species + the reference's step sequence."""

    def __init__(self, backend=None, host=None, n_pipeline=1):
        self.be = backend if backend is not None else b200_backend()
        self.host = host if host is not None else B200Host()
        self.n_pipeline = n_pipeline
        self.species = []          # definition order; the reference's list head is the LAST one
        self.num_comm_round = 3    # src/vpic/vpic.cc:79
        # divergence cleaning / shared-face synchronisation cadence (vpic.cc:72-78: off by default)
        self.clean_div_e_interval = self.clean_div_b_interval = self.sync_shared_interval = 0
        self.num_div_e_round = self.num_div_b_round = 2
        self.last_div_e_err = self.last_div_b_err = self.last_sync_err = None
        self.g = self.fa = self.ia = self.aa = None
        self.pbc_list = None       # particle_bc_t list head (user handlers), src/vpic/vpic.h:561-571
        self._pbc_keep = []

    # ---- setup (deck-like helpers, src/vpic/vpic.h:326-653) ----
    def define_grid(self, kind, lo, hi, n, topo=(1, 1, 1), dt=None, cvac=1.0, eps0=1.0, pbc=-2):
        g = self.host.new_grid()
        gs = self.grid_struct(g)
        gs.cvac, gs.eps0 = cvac, eps0
        gs.dt = dt if dt is not None else 0.0
        if kind == "periodic":
            self.host.partition_periodic_box(g, lo, hi, n, topo)
        elif kind == "metal":
            self.host.partition_metal_box(g, lo, hi, n, topo)
        elif kind == "absorbing":
            self.host.partition_absorbing_box(g, lo, hi, n, topo, pbc)
        else:
            raise ValueError(kind)
        self.g = g
        return g

    @staticmethod
    def grid_struct(g):
        return C.cast(g, C.POINTER(abi.Grid)).contents

    def neighbor_table(self):
        """g->neighbor as an (nv, 6) int64 view: the table set_region_bc edits
        (deck/wrapper.h:255-310)."""
        import numpy as np
        g = self.grid_struct(self.g)
        return np.ctypeslib.as_array(g.neighbor, shape=(g.nv, 6))

    def define_fields(self, damp=0.0, eps=1.0, mu=1.0, sigma=0.0, materials=()):
        self.m_list = self.host.material_list(eps, mu, sigma, materials)
        self.fa = self.host.field_array(self.g, self.m_list, damp)
        self.ia = self.host.interpolator_array(self.g)
        self.aa = self.host.accumulator_array(self.g, self.n_pipeline)

    def define_species(self, name, q, m, max_np, max_nm=None, sort_interval=0, sort_out_of_place=1, host_arrays=True):
        if max_nm is None:
            max_nm = max(16, int(0.05 * max_np))
        sp = self.host.species(name, q, m, max_np, max_nm, sort_interval, sort_out_of_place, self.g, host_arrays)
        s = self.sp_struct(sp)
        s.id = len(self.species)
        # append_species pushes on the front of the list (species_advance.c:96-98)
        s.next = C.cast(self.species[-1], C.POINTER(abi.Species)) if self.species else None
        self.species.append(sp)
        return sp

    def define_particle_bc(self, interact, params=None):
        """append_particle_bc (src/boundary/boundary.c:66-76): `interact` is the address of a host
        function with the particle_bc_func_t signature; returns the id (-3, -4, ...) to put on faces
        with set_pbc / define_grid(pbc=...)."""
        pbc = abi.ParticleBC()
        pbc.params, pbc.interact, pbc.delete_pbc = params, interact, None
        n = 0 if self.pbc_list is None else -C.cast(self.pbc_list, C.POINTER(abi.ParticleBC)).contents.id - 2
        pbc.id = -3 - n
        pbc.next = C.cast(self.pbc_list, C.POINTER(abi.ParticleBC)) if self.pbc_list else None
        self._pbc_keep.append(pbc)
        self.pbc_list = C.addressof(pbc)
        return pbc.id

    @staticmethod
    def sp_struct(sp):
        return C.cast(sp, C.POINTER(abi.Species)).contents

    @property
    def sp_list_head(self):
        return self.species[-1] if self.species else None

    def species_in_list_order(self):
        return list(reversed(self.species))

    # ---- array views ----
    def particles(self, sp):
        s = self.sp_struct(sp)
        return abi.as_array(s.p, s.np, abi.particle_dtype)

    def set_particles(self, sp, arr):
        s = self.sp_struct(sp)
        n = len(arr)
        if n > s.max_np:
            raise ValueError("too many particles")
        abi.as_array(s.p, s.max_np, abi.particle_dtype)[:n] = arr
        s.np = n
        s.nm = 0

    def fields(self):
        fa = C.cast(self.fa, C.POINTER(abi.FieldArray)).contents
        return abi.as_array(fa.f, self.grid_struct(self.g).nv, abi.field_dtype)

    def hydro(self):
        """The hydro array (created on first use) as a numpy view."""
        if getattr(self, "ha", None) is None:
            self.ha = self.host.hydro_array(self.g)
        ha = C.cast(self.ha, C.POINTER(abi.HydroArray)).contents
        return abi.as_array(ha.h, self.grid_struct(self.g).nv, abi.hydro_dtype)

    def hydro_moments(self, sp):
        """What dump_hydro computes for one species (src/vpic/dump.cc:227-229):
        clear, accumulate, synchronize."""
        self.hydro()
        self.be.clear_hydro_array(self.ha)
        self.be.accumulate_hydro_p(self.ha, sp, self.ia)
        self.be.synchronize_hydro_array(self.ha)
        return self.hydro()

    def interpolators(self):
        ia = C.cast(self.ia, C.POINTER(abi.InterpolatorArray)).contents
        return abi.as_array(ia.i, self.grid_struct(self.g).nv, abi.interp_dtype)

    def accumulators(self):
        aa = C.cast(self.aa, C.POINTER(abi.AccumulatorArray)).contents
        return abi.as_array(aa.a, (aa.n_pipeline + 1) * aa.stride, np.dtype(("f4", (12,))))

    def partition(self, sp):
        s = self.sp_struct(sp)
        return abi.as_array(s.partition, self.grid_struct(self.g).nv + 1, np.int32)

    def movers(self, sp):
        s = self.sp_struct(sp)
        return abi.as_array(s.pm, s.nm, abi.mover_dtype)

    # ---- initialize(): src/vpic/initialize.cc:14-54 ----
    def initialize(self, uncenter=True, full=False):
        """full=True runs the reference's whole post-deck sequence (shared-face
        check, div B clean, tca and rhob initialisation, div E clean); the
        default is its tail, which is all a deck with consistent fields needs."""
        be = self.be
        if full:
            fa = self.fa
            self.init_report = rep = {}
            rep["sync0"] = be.synchronize_tang_e_norm_b(fa)            # :16
            be.compute_div_b_err(fa)                                    # :20-23
            rep["div_b"] = be.compute_rms_div_b_err(fa)
            be.clean_div_b(fa)
            be.compute_curl_b(fa)                                       # :28
            be.clear_rhof(fa)                                           # :31-34
            for sp in self.species_in_list_order():
                be.accumulate_rho_p(fa, sp)
            be.synchronize_rho(fa)
            be.compute_rhob(fa)
            be.compute_div_e_err(fa)                                    # :40-43
            rep["div_e"] = be.compute_rms_div_e_err(fa)
            be.clean_div_e(fa)
            rep["sync1"] = be.synchronize_tang_e_norm_b(fa)            # :46
        be.load_interpolator_array(self.ia, self.fa)
        if uncenter:
            for sp in self.species_in_list_order():
                be.uncenter_p(sp, self.ia)

    # ---- advance(): src/vpic/advance.cc:15-208 ----
    def advance(self):
        be, g = self.be, self.grid_struct(self.g)
        for sp in self.species_in_list_order():                       # :25-29
            s = self.sp_struct(sp)
            if s.sort_interval > 0 and (g.step % s.sort_interval) == 0:
                be.sort_p(sp)
        be.clear_accumulator_array(self.aa)                           # :36-37
        for sp in self.species_in_list_order():                       # :49-50
            be.advance_p(sp, self.aa, self.ia)
        be.reduce_accumulator_array(self.aa)                          # :65-66
        for _ in range(self.num_comm_round):                          # :73-77
            be.boundary_p(self.pbc_list, self.sp_list_head, self.fa, self.aa)
        if any(self.sp_struct(sp).nm for sp in self.species):         # :78-101 (drop unprocessed movers into rhob)
            self.drop_unprocessed_movers()
        be.clear_jf(self.fa)                                          # :107-110
        be.unload_accumulator_array(self.fa, self.aa)
        be.synchronize_jf(self.fa)
        be.advance_b(self.fa, 0.5)                                    # :123
        be.advance_e(self.fa, 1.0)                                    # :127
        be.advance_b(self.fa, 0.5)                                    # :137
        step = g.step
        if self.clean_div_e_interval > 0 and step % self.clean_div_e_interval == 0:      # :141-157
            be.clear_rhof(self.fa)
            for sp in self.species_in_list_order():
                be.accumulate_rho_p(self.fa, sp)
            be.synchronize_rho(self.fa)
            for rnd in range(self.num_div_e_round):
                be.compute_div_e_err(self.fa)
                if rnd == 0 or rnd == self.num_div_e_round - 1:
                    self.last_div_e_err = be.compute_rms_div_e_err(self.fa)
                be.clean_div_e(self.fa)
        if self.clean_div_b_interval > 0 and step % self.clean_div_b_interval == 0:      # :161-171
            for rnd in range(self.num_div_b_round):
                be.compute_div_b_err(self.fa)
                if rnd == 0 or rnd == self.num_div_b_round - 1:
                    self.last_div_b_err = be.compute_rms_div_b_err(self.fa)
                be.clean_div_b(self.fa)
        if self.sync_shared_interval > 0 and step % self.sync_shared_interval == 0:      # :175-179
            self.last_sync_err = be.synchronize_tang_e_norm_b(self.fa)
        be.load_interpolator_array(self.ia, self.fa)                  # :185
        g.step += 1                                                   # :187

    def drop_unprocessed_movers(self):
        """src/vpic/advance.cc:78-101: movers that num_comm_round boundary_p calls did not
        finish are dropped -- voxel restored, charge to rhob, back-fill, last mover first."""
        if self.be.kind == "b200":
            self.be.L.vpic_b200_drop_unprocessed_movers(self.sp_list_head, self.fa)
            return
        fa = C.cast(self.fa, C.POINTER(abi.FieldArray)).contents
        for sp in self.species_in_list_order():
            s = self.sp_struct(sp)
            p, pm = abi.as_array(s.p, s.max_np, abi.particle_dtype), self.movers(sp)
            for k in range(s.nm - 1, -1, -1):
                i = int(pm["i"][k])
                p["i"][i] >>= 3
                self.be.accumulate_rhob(fa.f, C.c_void_p(C.cast(s.p, C.c_void_p).value + 32 * i), self.g, s.q)
                p[i] = p[s.np - 1]
                s.np -= 1
            s.nm = 0

    # ---- diagnostics: dump_energies (src/vpic/dump.cc:37-77) ----
    def energies(self):
        en = self.be.energy_f6(self.fa)
        for sp in self.species_in_list_order():
            en.append(self.be.energy_p(sp, self.ia))
        return en
