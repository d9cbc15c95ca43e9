"""ctypes binding of include/ls1g.h.  `Container` mirrors the reference's ParticleContainer + Integrator
operators for the hot path (names follow src/particleContainer/ParticleContainer.h and
src/integrators/Integrator.h): update, exchange (balanceAndExchange), traverse_cells, delete_outer_particles,
event_new_timestep, event_forces_calculated."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

HALO_FLAG = 0x80000000

EXPORTS = [
    "ls1g_last_error", "ls1g_version", "ls1g_create", "ls1g_destroy", "ls1g_set_components", "ls1g_get_grid",
    "ls1g_cell_index_of_points_host", "ls1g_cell_index_of_points", "ls1g_debug_cell_pairs", "ls1g_upload",
    "ls1g_upload_forces", "ls1g_stream", "ls1g_count", "ls1g_download", "ls1g_update", "ls1g_exchange", "ls1g_traverse", "ls1g_delete_outer",
    "ls1g_integrate_pre", "ls1g_integrate_post", "ls1g_scale", "ls1g_selective_thermostat", "ls1g_run", "ls1g_step", "ls1g_host_alloc", "ls1g_host_free", "ls1g_decomp_box", "ls1g_decomp_grid",
    "ls1g_comm_unique_id", "ls1g_comm_init", "ls1g_allreduce_sum", "ls1g_comm_init_local", "ls1g_exchange_group",
    "ls1g_timers", "ls1g_measure_fp64_peak", "ls1g_last_pair_count", "ls1g_flop_count", "ls1g_rdf", "ls1g_reserve", "ls1g_download_range", "ls1g_slot_count",
    "ls1g_run_group", "ls1g_step_group", "ls1g_list_stats", "ls1g_cell_index_of_points_noalloc", "ls1g_device_count",
    "ls1g_record_size", "ls1g_upload_records", "ls1g_download_records",
]

FIELDS = dict(id=1, cid=2, r=4, v=8, q=16, L=32, F=64, M=128, Vi=256, cell=512)
T_FORCE, T_SORT, T_INTEGRATE, T_EXCHANGE = 0, 1, 2, 3


class Ls1gError(RuntimeError):
    pass


class _Config(C.Structure):
    _fields_ = [("device", C.c_int), ("box_min", C.c_double * 3), ("box_max", C.c_double * 3),
                ("global_length", C.c_double * 3), ("cutoff", C.c_double), ("lj_cutoff", C.c_double),
                ("eps_rf", C.c_double), ("grid_mode", C.c_int), ("rank_grid", C.c_int * 3),
                ("rank_coords", C.c_int * 3), ("capacity", C.c_uint64)]


class _Component(C.Structure):
    _fields_ = [("mass", C.c_double), ("inertia", C.c_double * 3), ("rot_dof", C.c_int),
                ("n_lj", C.c_int), ("n_charge", C.c_int), ("n_dipole", C.c_int), ("n_quadrupole", C.c_int),
                ("lj", C.c_void_p), ("charge", C.c_void_p), ("dipole", C.c_void_p), ("quadrupole", C.c_void_p)]


class _Fields(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("id", "cid", "r", "v", "q", "L", "F", "M", "Vi", "cell")]


class _FlopCounts(C.Structure):
    _fields_ = [("molecule_distances", C.c_double), ("kernel_calls", C.c_double * 7), ("macro_calls", C.c_double * 7), ("flops", C.c_double)]


class _RdfParams(C.Structure):
    _fields_ = [("bins", C.c_uint32), ("angular_bins", C.c_uint32), ("interval_length", C.c_double)]


class _StepRecord(C.Structure):
    _fields_ = [("upot", C.c_double), ("virial", C.c_double), ("summv2", C.c_double), ("sumIw2", C.c_double),
                ("n", C.c_uint64), ("rot_dof", C.c_uint64)]


def lib_path():
    # LS1G_LIB: another build of the same library (A/B timing of two builds on one box); default = the in-tree build
    return os.environ.get("LS1G_LIB") or os.path.join(_HERE, "libls1g.so")


def build(force=False):
    """compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)"""
    args = ["make", "-s", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    if force:
        subprocess.check_call(args + ["clean"])
    subprocess.check_call(args)


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(lib_path()):
            raise Ls1gError("libls1g.so is not built (run __graft_entry__.build()); there is no CPU fallback")
        _LIB = C.CDLL(lib_path())
        _LIB.ls1g_last_error.restype = C.c_char_p
    return _LIB


def _chk(rc):
    if rc != 0:
        raise Ls1gError(lib().ls1g_last_error().decode())


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def decomp_grid(n_ranks, rank):
    g = (C.c_int * 3)()
    co = (C.c_int * 3)()
    _chk(lib().ls1g_decomp_grid(int(n_ranks), int(rank), g, co))
    return list(g), list(co)


def decomp_box(grid, coords, global_length):
    bmin = (C.c_double * 3)()
    bmax = (C.c_double * 3)()
    _chk(lib().ls1g_decomp_box((C.c_int * 3)(*grid), (C.c_int * 3)(*coords), (C.c_double * 3)(*global_length), bmin, bmax))
    return list(bmin), list(bmax)


def device_count():
    """ls1g_device_count: CUDA devices this process can use (0 on a host without a usable GPU)"""
    n = C.c_int(0)
    _chk(lib().ls1g_device_count(C.byref(n)))
    return n.value


def fp64_peak_tflops(device=0):
    t = C.c_double()
    _chk(lib().ls1g_measure_fp64_peak(int(device), C.byref(t)))
    return t.value


def comm_init_local(containers):
    arr = (C.c_void_p * len(containers))(*[c._ctx for c in containers])
    _chk(lib().ls1g_comm_init_local(arr, len(containers)))


def exchange_group(containers):
    arr = (C.c_void_p * len(containers))(*[c._ctx for c in containers])
    _chk(lib().ls1g_exchange_group(arr, len(containers)))


def run_group(containers, n_steps, dt, target_temperature=0.0):
    """ls1g_run for all ranks of a local group in lockstep; returns [n_steps][6] (upot virial summv2 sumIw2 N rotDOF), summed over the ranks"""
    arr = (C.c_void_p * len(containers))(*[c._ctx for c in containers])
    rec = (_StepRecord * max(n_steps, 1))()
    _chk(lib().ls1g_run_group(arr, len(containers), int(n_steps), C.c_double(dt), C.c_double(target_temperature), rec))
    return np.array([[x.upot, x.virial, x.summv2, x.sumIw2, x.n, x.rot_dof] for x in rec[:n_steps]], dtype=np.float64).reshape(n_steps, 6)


def step_group(containers, dt, beta_trans=1.0, beta_rot=1.0, vi=True):
    """ls1g_step for all ranks of a local group; returns the step record summed over the ranks"""
    arr = (C.c_void_p * len(containers))(*[c._ctx for c in containers])
    rec = _StepRecord()
    _chk(lib().ls1g_step_group(arr, len(containers), C.c_double(dt), C.c_double(beta_trans), C.c_double(beta_rot), int(bool(vi)), C.byref(rec)))
    return np.array([rec.upot, rec.virial, rec.summv2, rec.sumIw2, rec.n, rec.rot_dof], dtype=np.float64)


def comm_unique_id():
    """NCCL bootstrap id (rank 0 creates it; distribute the 128 bytes to all ranks)."""
    buf = (C.c_ubyte * 128)()
    _chk(lib().ls1g_comm_unique_id(buf))
    return bytes(buf)


def components_from_tables(counts, scal, lj, ch, di, qu):
    """Flattened component tables (layout of the reference dump / ls1g_component rows) -> list of dicts.
    counts [nc][>=5]: nLJ nCharge nDipole nQuadrupole rotDOF; scal [nc][4]: m I11 I22 I33."""
    counts = np.asarray(counts, dtype=np.int64).reshape(len(np.atleast_2d(counts)), -1)
    scal = np.asarray(scal, dtype=np.float64).reshape(-1, 4)
    lj = np.asarray(lj, dtype=np.float64).reshape(-1, 7)
    ch = np.asarray(ch, dtype=np.float64).reshape(-1, 5)
    di = np.asarray(di, dtype=np.float64).reshape(-1, 7)
    qu = np.asarray(qu, dtype=np.float64).reshape(-1, 7)
    out, o = [], [0, 0, 0, 0]
    for k in range(len(counts)):
        n = [int(x) for x in counts[k, :4]]
        out.append(dict(mass=float(scal[k, 0]), inertia=[float(x) for x in scal[k, 1:4]], rot_dof=int(counts[k, 4]),
                        lj=lj[o[0]:o[0] + n[0]].copy(), charge=ch[o[1]:o[1] + n[1]].copy(),
                        dipole=di[o[2]:o[2] + n[2]].copy(), quadrupole=qu[o[3]:o[3] + n[3]].copy()))
        o = [o[j] + n[j] for j in range(4)]
    return out


def lj_component(eps=1.0, sigma=1.0, mass=1.0, shift6=0.0):
    """single-centre Lennard-Jones component (configs A, B, E)"""
    return dict(mass=mass, inertia=[0.0, 0.0, 0.0], rot_dof=0, lj=np.array([[0, 0, 0, mass, eps, sigma, shift6]], dtype=np.float64),
                charge=np.zeros((0, 5)), dipole=np.zeros((0, 7)), quadrupole=np.zeros((0, 7)))


class Container:
    """GPU-resident linked-cell particle container + leapfrog integrator of ONE rank.

    Mirrors, for the hot path only, ParticleContainer (src/particleContainer/ParticleContainer.h:85-269),
    Integrator (src/integrators/Integrator.h:33-90) and DomainDecompBase::balanceAndExchange; every method is a
    thin call into the C ABI.  Arrays are numpy, host memory."""

    def __init__(self, box_max, cutoff, box_min=(0.0, 0.0, 0.0), lj_cutoff=None, eps_rf=1e10, grid_mode=0,
                 global_length=None, rank_grid=(1, 1, 1), rank_coords=(0, 0, 0), device=0, capacity=0):
        cfg = _Config()
        cfg.device = int(device)
        cfg.box_min = (C.c_double * 3)(*[float(x) for x in box_min])
        cfg.box_max = (C.c_double * 3)(*[float(x) for x in box_max])
        gl = box_max if global_length is None else global_length
        cfg.global_length = (C.c_double * 3)(*[float(x) for x in gl])
        cfg.cutoff = float(cutoff)
        cfg.lj_cutoff = float(cutoff if lj_cutoff is None else lj_cutoff)
        cfg.eps_rf = float(eps_rf)
        cfg.grid_mode = int(grid_mode)
        cfg.rank_grid = (C.c_int * 3)(*[int(x) for x in rank_grid])
        cfg.rank_coords = (C.c_int * 3)(*[int(x) for x in rank_coords])
        cfg.capacity = int(capacity)
        self._ctx = C.c_void_p()
        self._keep = []
        _chk(lib().ls1g_create(C.byref(cfg), C.byref(self._ctx)))

    # -- life cycle ---------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx:
            lib().ls1g_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- parameters ---------------------------------------------------------------------------------
    def set_components(self, comps, mixcoeff=()):
        """comps: list of dicts (mass, inertia[3], rot_dof, lj[n][7], charge[n][5], dipole[n][7], quadrupole[n][7]);
        mixcoeff: (xi, eta) per component pair i<j (Domain::getmixcoeff)."""
        arr = (_Component * len(comps))()
        keep = []
        for k, c in enumerate(comps):
            a = arr[k]
            a.mass = float(c["mass"])
            a.inertia = (C.c_double * 3)(*[float(x) for x in c["inertia"]])
            a.rot_dof = int(c["rot_dof"])
            for name, fld, w in (("lj", "n_lj", 7), ("charge", "n_charge", 5), ("dipole", "n_dipole", 7), ("quadrupole", "n_quadrupole", 7)):
                t = np.ascontiguousarray(c.get(name, np.zeros((0, w))), dtype=np.float64).reshape(-1, w)
                keep.append(t)
                setattr(a, fld, len(t))
                setattr(a, name, _p(t) if len(t) else None)
        mix = np.ascontiguousarray(mixcoeff, dtype=np.float64).ravel()
        _chk(lib().ls1g_set_components(self._ctx, len(comps), arr, _p(mix) if len(mix) else None, int(len(mix))))
        self.n_components = len(comps)
        self.n_sites = [sum(len(np.reshape(c.get(k, ()), (-1,))) // w for k, w in (("lj", 7), ("charge", 5), ("dipole", 7), ("quadrupole", 7)))
                        for c in comps]

    # -- grid ---------------------------------------------------------------------------------------
    def get_grid(self):
        cpd = (C.c_int * 3)()
        cl = (C.c_double * 3)()
        _chk(lib().ls1g_get_grid(self._ctx, cpd, cl))
        return list(cpd), list(cl)

    def cell_index_of_points(self, pts, host=False):
        pts = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 3)
        out = np.empty(len(pts), dtype=np.uint32)
        f = lib().ls1g_cell_index_of_points_host if host else lib().ls1g_cell_index_of_points
        _chk(f(self._ctx, C.c_size_t(len(pts)), _p(pts), _p(out)))
        return out

    NOALLOC_MAX_POINTS = 65536      # LS1G_NOALLOC_MAX_POINTS

    def cell_index_of_points_noalloc(self, pts):
        """ls1g_cell_index_of_points_noalloc: same kernel, device staging in a fixed __device__ array (no cudaMalloc in the call)"""
        pts = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 3)
        out = np.empty(len(pts), dtype=np.uint32)
        _chk(lib().ls1g_cell_index_of_points_noalloc(self._ctx, C.c_size_t(len(pts)), _p(pts), _p(out)))
        return out

    def cell_pairs(self):
        n = C.c_size_t()
        _chk(lib().ls1g_debug_cell_pairs(self._ctx, None, C.c_size_t(0), C.byref(n)))
        out = np.empty((n.value, 3), dtype=np.uint32)
        _chk(lib().ls1g_debug_cell_pairs(self._ctx, _p(out), C.c_size_t(n.value), C.byref(n)))
        return out

    # -- storage ------------------------------------------------------------------------------------
    def add_particles(self, id, cid, r, v=None, q=None, L=None):
        """ParticleContainer::addParticles: replaces the content."""
        id = np.ascontiguousarray(id, dtype=np.uint64)
        n = len(id)
        cid = np.ascontiguousarray(cid, dtype=np.uint32)
        r = np.ascontiguousarray(r, dtype=np.float64).reshape(n, 3)
        v = None if v is None else np.ascontiguousarray(v, dtype=np.float64).reshape(n, 3)
        q = None if q is None else np.ascontiguousarray(q, dtype=np.float64).reshape(n, 4)
        L = None if L is None else np.ascontiguousarray(L, dtype=np.float64).reshape(n, 3)
        _chk(lib().ls1g_upload(self._ctx, C.c_size_t(n), _p(id), _p(cid), _p(r), _p(v), _p(q), _p(L)))

    RECORD_FORMATS = dict(ICRVQD=0, ICRV=1, IRV=2)      # LS1G_REC_*

    @staticmethod
    def record_size(fmt="ICRVQD"):
        n = C.c_size_t()
        _chk(lib().ls1g_record_size(Container.RECORD_FORMATS[fmt] if isinstance(fmt, str) else int(fmt), C.byref(n)))
        return n.value

    def add_records(self, records, fmt="ICRVQD"):
        """ls1g_upload_records: the packed records of a binary checkpoint (bytes / uint8 array / structured array, as they stand in the
        .dat file) replace the content; the device unpacks them and keeps those inside this context's bounding box
        (BinaryReader::readPhaseSpace).  Returns (molecules kept, highest id of all records)."""
        buf = np.frombuffer(records, dtype=np.uint8) if isinstance(records, (bytes, bytearray, memoryview)) else np.ascontiguousarray(records).view(np.uint8).reshape(-1)
        size = self.record_size(fmt)
        if buf.size % size:
            raise Ls1gError("add_records: %d bytes are not a whole number of %d-byte records" % (buf.size, size))
        kept, mx = C.c_uint64(), C.c_uint64()
        _chk(lib().ls1g_upload_records(self._ctx, C.c_size_t(buf.size // size), _p(buf) if buf.size else None, self.RECORD_FORMATS[fmt], C.byref(kept), C.byref(mx)))
        return kept.value, mx.value

    def download_records(self, out=None, scale_v=1.0, scale_D=1.0):
        """ls1g_download_records: the owned molecules as ICRVQD records (uint8 array [n][116]), cell by cell - the bytes
        DomainDecompBase::writeMoleculesToFile(binary) appends to the .dat file.  out: optional preallocated (e.g. pinned) uint8 buffer;
        scale_v, scale_D: factors applied to v and D on the way out (a thermostat scaling the device has not applied yet)"""
        n = C.c_size_t()
        _chk(lib().ls1g_download_records(self._ctx, C.c_double(1.0), C.c_double(1.0), C.c_size_t(0), None, C.byref(n)))
        if out is not None:
            if out.dtype != np.uint8 or out.size < n.value * 116 or not out.flags.c_contiguous:
                raise Ls1gError("download_records: out must be a contiguous uint8 buffer of at least %d bytes" % (n.value * 116))
            out = out.reshape(-1)[:n.value * 116].reshape(n.value, 116)
        else:
            out = np.empty((n.value, 116), dtype=np.uint8)
        if n.value:
            _chk(lib().ls1g_download_records(self._ctx, C.c_double(scale_v), C.c_double(scale_D), C.c_size_t(n.value), _p(out), C.byref(n)))
        return out

    def set_forces(self, F, M=None):
        """Molecule::setF/setM for the molecules of the directly preceding add_particles (same order)"""
        F = np.ascontiguousarray(F, dtype=np.float64).reshape(-1, 3)
        M = None if M is None else np.ascontiguousarray(M, dtype=np.float64).reshape(-1, 3)
        _chk(lib().ls1g_upload_forces(self._ctx, C.c_size_t(len(F)), _p(F), _p(M)))

    def stream(self):
        """cudaStream_t (as int) all work of this container is issued on"""
        st = C.c_void_p()
        _chk(lib().ls1g_stream(self._ctx, C.byref(st)))
        return st.value or 0

    def count(self):
        a, b = C.c_uint64(), C.c_uint64()
        _chk(lib().ls1g_count(self._ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    def get_number_of_particles(self):
        return self.count()[0]

    def download(self, fields=("id", "cid", "r", "v", "q", "L", "F", "M", "Vi", "cell"), include_halo=False, out=None):
        """molecules in cell-sorted order (ParticleIterator over ONLY_INNER_AND_BOUNDARY / ALL_CELLS).
        out: optional dict of preallocated (e.g. pinned) arrays to fill instead of allocating"""
        n = C.c_size_t()
        _chk(lib().ls1g_download(self._ctx, 0, int(bool(include_halo)), C.c_size_t(0), C.byref(_Fields()), C.byref(n)))      # count query
        cap = n.value
        shapes = dict(id=((cap,), np.uint64), cid=((cap,), np.uint32), r=((cap, 3), np.float64), v=((cap, 3), np.float64),
                      q=((cap, 4), np.float64), L=((cap, 3), np.float64), F=((cap, 3), np.float64), M=((cap, 3), np.float64),
                      Vi=((cap, 3), np.float64), cell=((cap,), np.uint32))
        given, out, f, mask = out, {}, _Fields(), 0
        for k in fields:
            if given is not None:
                a = given[k]
                if a.dtype != shapes[k][1] or a.size < int(np.prod(shapes[k][0])) or not a.flags.c_contiguous:
                    raise Ls1gError("download: out[%r] has the wrong dtype/size" % k)
                out[k] = a.reshape(-1)[:int(np.prod(shapes[k][0]))].reshape(shapes[k][0])
            else:
                out[k] = np.zeros(*shapes[k])
            setattr(f, k, _p(out[k]))
            mask |= FIELDS[k]
        n = C.c_size_t()
        _chk(lib().ls1g_download(self._ctx, mask, int(bool(include_halo)), C.c_size_t(cap), C.byref(f), C.byref(n)))
        return {k: a[:n.value] for k, a in out.items()}

    def download_slots(self, fields, first=0, count=None):
        """raw slots [first, first + count) of the cell-sorted storage (ls1g_download_range): halo copies included, flagged with
        HALO_FLAG in 'cell'"""
        ns = C.c_uint64()
        _chk(lib().ls1g_slot_count(self._ctx, C.byref(ns)))
        count = ns.value - first if count is None else count
        widths = dict(id=((count,), np.uint64), cid=((count,), np.uint32), r=((count, 3), np.float64), v=((count, 3), np.float64),
                      q=((count, 4), np.float64), L=((count, 3), np.float64), F=((count, 3), np.float64), M=((count, 3), np.float64),
                      Vi=((count, 3), np.float64), cell=((count,), np.uint32))
        out, f, mask = {}, _Fields(), 0
        for k in fields:
            out[k] = np.zeros(*widths[k])
            setattr(f, k, _p(out[k]))
            mask |= FIELDS[k]
        _chk(lib().ls1g_download_range(self._ctx, mask, C.c_size_t(first), C.c_size_t(count), C.byref(f)))
        return out

    # -- per-step operators ---------------------------------------------------------------------------
    def update(self):
        _chk(lib().ls1g_update(self._ctx))

    def exchange(self):
        """DomainDecompBase::balanceAndExchange: wrap/send leaving molecules, halo copies, re-sort"""
        _chk(lib().ls1g_exchange(self._ctx))

    def traverse_cells(self):
        """force pass + calcFM; returns (upot, virial) as handed to Domain::setLocalUpot/Virial"""
        u, w = C.c_double(), C.c_double()
        _chk(lib().ls1g_traverse(self._ctx, C.byref(u), C.byref(w)))
        return u.value, w.value

    def delete_outer_particles(self):
        _chk(lib().ls1g_delete_outer(self._ctx))

    def event_new_timestep(self, dt):
        _chk(lib().ls1g_integrate_pre(self._ctx, C.c_double(dt)))

    def event_forces_calculated(self, dt):
        """returns (summv2, sumIw2, N, rotDOF) (Leapfrog.cpp:146-150)"""
        a, b = C.c_double(), C.c_double()
        n, r = C.c_uint64(), C.c_uint64()
        _chk(lib().ls1g_integrate_post(self._ctx, C.c_double(0.5 * dt), C.byref(a), C.byref(b), C.byref(n), C.byref(r)))
        return a.value, b.value, n.value, r.value

    def scale(self, beta_trans, beta_rot=1.0):
        _chk(lib().ls1g_scale(self._ctx, C.c_double(beta_trans), C.c_double(beta_rot)))

    def reserve(self, n):
        """allocate the device arrays for n owned molecules ahead of add_particles (set-up, like the constructor)"""
        _chk(lib().ls1g_reserve(self._ctx, C.c_size_t(int(n))))

    def flop_count(self):
        """the reference's FlopCounter over the current cells (halo copies in place): dict(molecule_distances, kernel[7], macro[7], flops)"""
        f = _FlopCounts()
        _chk(lib().ls1g_flop_count(self._ctx, C.byref(f)))
        return dict(molecule_distances=f.molecule_distances, kernel=np.array(list(f.kernel_calls)), macro=np.array(list(f.macro_calls)), flops=f.flops)

    def rdf(self, bins, interval_length, angular_bins=1):
        """the reference's RDFCellProcessor + RDF counters over the current cells (halo copies in place): dict(mol[nc(nc+1)/2, bins],
        site[S, S, bins] over the global site rows, ardf[nc, nc, bins*angular_bins]) of this one traversal"""
        nc, S = self.n_components, int(sum(self.n_sites))
        mol = np.zeros((nc * (nc + 1) // 2, bins), dtype=np.uint64)
        site = np.zeros((S, S, bins), dtype=np.uint64)
        ardf = np.zeros((nc, nc, bins * max(angular_bins, 1)), dtype=np.uint64)
        p = _RdfParams(int(bins), int(angular_bins), float(interval_length))
        _chk(lib().ls1g_rdf(self._ctx, C.byref(p), _p(mol), _p(site), _p(ardf) if angular_bins > 1 else None))
        return dict(mol=mol, site=site, ardf=ardf if angular_bins > 1 else ardf[:, :, :0])

    def selective_thermostat(self, limit_trans_energy, temperature):
        """Domain::calculateGlobalValues explosion heuristics (Domain.cpp:254-295): cap U_trans / U_rot per molecule"""
        _chk(lib().ls1g_selective_thermostat(self._ctx, C.c_double(limit_trans_energy), C.c_double(temperature)))

    def run(self, n_steps, dt, target_temperature=0.0, records=True):
        """device-resident time loop; returns array [n_steps][6]: upot virial summv2 sumIw2 N rotDOF"""
        rec = (_StepRecord * max(n_steps, 1))() if records else None
        _chk(lib().ls1g_run(self._ctx, int(n_steps), C.c_double(dt), C.c_double(target_temperature), rec))
        if not records:
            return None
        return np.array([[x.upot, x.virial, x.summv2, x.sumIw2, x.n, x.rot_dof] for x in rec[:n_steps]], dtype=np.float64).reshape(n_steps, 6)

    def step(self, dt, beta_trans=1.0, beta_rot=1.0, vi=True):
        """one host-driven time step (ls1g_step): previous step's thermostat betas fused into the pre-kick; returns the step record"""
        rec = _StepRecord()
        _chk(lib().ls1g_step(self._ctx, C.c_double(dt), C.c_double(beta_trans), C.c_double(beta_rot), int(bool(vi)), C.byref(rec)))
        return np.array([rec.upot, rec.virial, rec.summv2, rec.sumIw2, rec.n, rec.rot_dof], dtype=np.float64)

    # -- multi-GPU ------------------------------------------------------------------------------------
    def comm_init(self, n_ranks, rank, unique_id):
        buf = (C.c_ubyte * 128).from_buffer_copy(unique_id)
        _chk(lib().ls1g_comm_init(self._ctx, int(n_ranks), int(rank), buf))

    def allreduce_sum(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64).copy()
        _chk(lib().ls1g_allreduce_sum(self._ctx, _p(v), int(len(v))))
        return v

    # -- measurement ----------------------------------------------------------------------------------
    def timers(self, reset=False):
        ms = (C.c_double * 5)()
        ln = (C.c_uint64 * 5)()
        _chk(lib().ls1g_timers(self._ctx, ms, ln, int(reset)))
        names = ("force", "sort", "integrate", "exchange", "aux")
        return {k: dict(ms=ms[i], launches=int(ln[i])) for i, k in enumerate(names)}

    def list_stats(self):
        """neighbour-list bookkeeping: dict(rebuilds, rows_per_block, entries_per_molecule, skin)"""
        a, b, c, d = C.c_uint64(), C.c_int(), C.c_int(), C.c_double()
        _chk(lib().ls1g_list_stats(self._ctx, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return dict(rebuilds=a.value, rows_per_block=b.value, entries_per_molecule=c.value, skin=d.value)

    def last_pair_count(self):
        n = C.c_uint64()
        _chk(lib().ls1g_last_pair_count(self._ctx, C.byref(n)))
        return n.value
