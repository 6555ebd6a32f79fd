"""Single-resolution incompressible Navier-Stokes stepper on the fused sm_100a kernel.

Mirrors ``IncompressibleNavierStokesStepper`` of the reference (xlb/operator/stepper/nse_stepper.py:41-282):
same constructor, ``prepare_fields(initializer=None) -> (f_0, f_1, bc_mask, missing_mask)`` and
``stepper(f_0, f_1, bc_mask, missing_mask, omega, timestep) -> (f_0, f_1)`` with the result in ``f_1`` and the
buffer swap left to the caller (examples/performance/mlups_3d.py:227-229).  One call = one launch of
``xlb_step`` on the current torch CUDA stream; there is no host synchronisation inside.
"""

import os

import numpy as np
import torch

from ... import _lib
from ...compute_backend import ComputeBackend
from ...default_config import DefaultConfig, current_device
from ...helper.check_boundary_overlaps import check_bc_overlaps
from ...helper.nse_fields import create_nse_fields
from ..boundary_condition.boundary_condition import ImplementationStep
from ..boundary_masker import IndicesBoundaryMasker
from ..collision import BGK, KBC, ForcedCollision, SmagorinskyLESBGK
from ..equilibrium import QuadraticEquilibrium
from ..macroscopic import Macroscopic
from ..operator import Operator, byref, check, lib, stream_ptr
from ..stream import Stream
from .stepper import Stepper


class IncompressibleNavierStokesStepper(Stepper):
    def __init__(
        self,
        grid,
        boundary_conditions=[],
        collision_type="BGK",
        streaming_scheme="pull",
        forcing_scheme="exact_difference",
        force_vector=None,
        backend_config={},
    ):
        super().__init__(grid, boundary_conditions)
        self.backend_config = backend_config
        if collision_type == "BGK":
            self.collision = BGK(self.velocity_set, self.precision_policy, self.compute_backend)
        elif collision_type == "KBC":
            self.collision = KBC(self.velocity_set, self.precision_policy, self.compute_backend)
        elif collision_type == "SmagorinskyLESBGK":
            self.collision = SmagorinskyLESBGK(self.velocity_set, self.precision_policy, self.compute_backend)
        else:
            raise NotImplementedError(f"collision_type={collision_type!r}: expected 'BGK', 'KBC' or 'SmagorinskyLESBGK' (nse_stepper.py:79-84)")
        if force_vector is not None:  # nse_stepper.py:86-87
            if not hasattr(force_vector, "shape"):
                force_vector = np.asarray(force_vector, dtype=np.float64)
            self.collision = ForcedCollision(collision_operator=self.collision, forcing_scheme=forcing_scheme, force_vector=force_vector)
        self._collision_params = self.collision.abi_params() or _lib.XlbCollisionParams()
        for bc in self.boundary_conditions:  # which BC kinds the kernel has to be able to handle (0 would mean "any")
            self._collision_params.bc_kind_mask |= 1 << int(bc.abi_kind)
        # backend_config["in_place"]: one population buffer instead of two (the in-place AA pattern, csrc/xlb_step_aa.cuh):
        # prepare_fields() returns the SAME tensor as f_0 and f_1, the usual loop ``f_0, f_1 = stepper(f_0, f_1, ...); swap``
        # runs unchanged, and after every even number of steps the buffer holds exactly what f_0 would hold in the two-buffer
        # scheme (after an odd number it is in the pattern's intermediate layout: read results on even step counts).
        self.in_place = bool(self.backend_config.get("in_place", False)) if self.backend_config else False
        self.aa_parity = 0  # 0: the next in-place step is a P step (buffer in the normal layout), 1: an L step
        if self.in_place:
            bad = [bc.__class__.__name__ for bc in self.boundary_conditions if int(bc.abi_kind) not in (_lib.BC_FULLWAY, _lib.BC_EQUILIBRIUM)]
            if bad:
                raise NotImplementedError(
                    f"in_place stepping supports FullwayBounceBackBC and EquilibriumBC only; {sorted(set(bad))} read f_pre or a "
                    "neighbour cell and need the second buffer")
            if force_vector is not None:
                raise NotImplementedError("in_place stepping: the body-force collisions keep the two-buffer kernels")
            if getattr(grid, "halo", 0):
                raise NotImplementedError("in_place stepping runs on a single block (no slab decomposition)")
        self.streaming_scheme = streaming_scheme
        if streaming_scheme != "pull":
            raise NotImplementedError(f"Unknown or unimplemented streaming scheme for backend: {self.compute_backend}")  # nse_stepper.py:91-92
        self.stream = Stream(self.velocity_set, self.precision_policy, self.compute_backend)
        self.equilibrium = QuadraticEquilibrium(self.velocity_set, self.precision_policy, self.compute_backend)
        self.macroscopic = Macroscopic(self.velocity_set, self.precision_policy, self.compute_backend)
        self.kernel_variant = int(self.backend_config.get("kernel_variant", _lib.VARIANT_AUTO)) if self.backend_config else _lib.VARIANT_AUTO
        self._plan = None  # (key, domain struct, bc table tensor, n_bc)
        # BC cell list (XlbCollisionParams.bc_cells): (key, sorted cell indices on the device, per-plane offsets on the host)
        self._cells = None
        self._mbits = None  # (key, missing_mask packed to one uint32 per cell)
        # backend_config["bc_cell_list"]: True / False force it; "auto" (default) uses it when the block holds enough such
        # cells for the second launch to pay for itself (B200: ~1e5 cells; below that the in-kernel path is ahead,
        # profiles/r1q_*: 384x128x128 sphere 22.4k -> 21.6k MLUPS with the list, 192x1536x1536 wind tunnel 21.5k -> 23.9k)
        self._use_cell_list = self.backend_config.get("bc_cell_list", "auto") if self.backend_config else "auto"
        env = os.environ.get("XLB_B200_BC_CELL_LIST")  # experiments: "1" / "0" override the default for steppers that do not set it
        if env in ("0", "1") and not (self.backend_config and "bc_cell_list" in self.backend_config):
            self._use_cell_list = env == "1"

    # ------------------------------------------------------------------ setup (nse_stepper.py:99-205)
    def prepare_fields(self, initializer=None):
        if self.in_place:  # one population buffer: never allocate the second one, not even transiently
            from ...precision_policy import Precision

            f_0 = self.grid.create_field(cardinality=self.velocity_set.q, dtype=self.precision_policy.store_precision)
            f_1 = f_0
            missing_mask = self.grid.create_field(cardinality=self.velocity_set.q, dtype=Precision.BOOL)
            bc_mask = self.grid.create_field(cardinality=1, dtype=Precision.UINT8)
            self.aa_parity = 0
        else:
            _, f_0, f_1, missing_mask, bc_mask = create_nse_fields(
                grid=self.grid, velocity_set=self.velocity_set, compute_backend=self.compute_backend, precision_policy=self.precision_policy
            )
        self._collision_params.halfway_fraction = self._halfway_fraction(bc_mask)
        bc_mask, missing_mask = self._process_boundary_conditions(self.boundary_conditions, bc_mask, missing_mask)
        if initializer is not None:
            f_0 = initializer(bc_mask, f_0)
        else:
            from ...helper.initializers import initialize_eq

            if f_0.numel() >= (1 << 28):
                # rho = 1, u = 0 gives feq_l = w_l exactly; skip the rho/u temporaries on very large grids
                w = self.velocity_set.w.astype(self.precision_policy.compute_precision.np_dtype)
                for l in range(self.velocity_set.q):
                    f_0[l].fill_(float(w[l]))
            else:
                f_0 = initialize_eq(f_0, self.grid, self.velocity_set, self.precision_policy, self.compute_backend)
        if self.boundary_conditions and not self.in_place:
            self._bc_cell_list(f_0, bc_mask)  # built here so that the step loop itself never synchronises
            self._missing_bits(f_0, missing_mask)
        return f_0, (f_0 if self.in_place else f_1), bc_mask, missing_mask

    def _halfway_fraction(self, bc_mask):
        """Share of this block's cells listed by HalfwayBounceBackBC objects (obstacle interiors included): a hint for the
        kernel-variant choice of ``xlb_step_ex`` (``XlbCollisionParams.halfway_fraction``)."""
        n = 0
        x0 = getattr(self.grid, "x_offset", 0)
        nx_local = self.grid.shape[0]
        for bc in self.boundary_conditions:
            idx = getattr(bc, "indices", None)
            if idx is None or int(bc.abi_kind) != _lib.BC_HALFWAY:
                continue
            xs = np.asarray(idx[0] if not isinstance(idx, torch.Tensor) else idx[0].cpu())
            n += int(np.count_nonzero((xs >= x0) & (xs < x0 + nx_local))) if getattr(self.grid, "halo", 0) else int(xs.size)
        return n / max(1, bc_mask.numel())

    def _process_boundary_conditions(self, boundary_conditions, bc_mask, missing_mask):
        spent = [bc for bc in boundary_conditions if "indices" not in bc.__dict__ or "mesh_vertices" not in bc.__dict__]
        if spent:
            # the maskers consume bc.indices / bc.mesh_vertices (indices_boundary_masker.py:131, mesh_boundary_masker.py:203-204);
            # the reference fails with an AttributeError on the second prepare_fields() -- fail as loudly, never mask nothing
            raise AttributeError(
                f"{spent[0].__class__.__name__} (id {spent[0].id}) has neither indices nor mesh_vertices left: prepare_fields() was already "
                "called with these boundary-condition objects; build new BC objects (or clone the masks of the first call)"
            )
        check_bc_overlaps(boundary_conditions, self.velocity_set.d, self.compute_backend)
        with_indices = [bc for bc in boundary_conditions if getattr(bc, "indices", None) is not None]
        if with_indices:
            masker = IndicesBoundaryMasker(
                velocity_set=self.velocity_set, precision_policy=self.precision_policy, compute_backend=self.compute_backend, grid=self.grid
            )
            start = None
            gshape = None
            if getattr(self.grid, "halo", 0) or getattr(self.grid, "x_offset", 0):
                start = (self.grid.x_offset - self.grid.halo,) + (0,) * (self.velocity_set.d - 1)
                gshape = self.grid.global_shape
            bc_mask, missing_mask = masker(with_indices, bc_mask, missing_mask, start, gshape)
        # mesh-based boundary conditions (nse_stepper.py:165-203): voxelised on the device, after the index-based ones
        with_vertices = [bc for bc in boundary_conditions if getattr(bc, "mesh_vertices", None) is not None]
        if with_vertices:
            if self.velocity_set.d != 3:
                raise NotImplementedError("mesh-based boundary conditions need a 3-D velocity set (mesh_boundary_masker.py:36-39)")
            from ..boundary_masker import MeshMaskerRay, MeshVoxelizationMethod

            start = None
            gshape = None
            if getattr(self.grid, "halo", 0) or getattr(self.grid, "x_offset", 0):
                start = (self.grid.x_offset - self.grid.halo, 0, 0)
                gshape = self.grid.global_shape
            for bc in with_vertices:
                method = bc.voxelization_method if bc.voxelization_method is not None else MeshVoxelizationMethod("RAY")
                if method.id != MeshVoxelizationMethod("RAY").id:
                    raise NotImplementedError(f"voxelization method {method.name}: the CUDA backend implements RAY (ray.py) only")
                if bc.needs_mesh_distance:
                    raise NotImplementedError("boundary conditions that need the mesh distance (HybridBC) are outside the CUDA backend's scope")
                mesh_masker = MeshMaskerRay(velocity_set=self.velocity_set, precision_policy=self.precision_policy, compute_backend=self.compute_backend)
                _, bc_mask, missing_mask = mesh_masker(bc, None, bc_mask, missing_mask, start, gshape)
        return bc_mask, missing_mask

    # ------------------------------------------------------------------ BC table
    def build_bc_table(self, field_shape):
        """256 ``XlbBcEntry`` rows (row = bc id) on the device."""
        table = np.zeros(256, dtype=_lib.BC_ENTRY_DTYPE)
        n_bc = 0
        for bc in self.boundary_conditions:
            if not (1 <= bc.id <= 253):
                raise ValueError(f"boundary-condition id {bc.id} does not fit bc_mask (1..253)")
            # global coordinates of this block's cell 0 (slab grids carry ghost planes in front of their first owned plane)
            bc._block_origin = (getattr(self.grid, "x_offset", 0) - getattr(self.grid, "halo", 0),) + (0,) * (self.velocity_set.d - 1)
            bc.bc_table_row(table[bc.id], field_shape)
            n_bc = max(n_bc, bc.id + 1)
        dev = torch.from_numpy(table.view(np.uint8).copy()).to(current_device())
        return dev, n_bc

    def _get_plan(self, f_0):
        key = (tuple(f_0.shape), f_0.dtype, tuple(bc.id for bc in self.boundary_conditions))
        if self._plan is None or self._plan[0] != key:
            halo = getattr(self.grid, "halo", 0)
            # a 3-D slab has ghost planes instead of the periodic wrap in x; a D2Q9 slab runs as nx = 1 with ghost ROWS
            dom = self._domain(f_0, x_periodic=0 if (halo and self.velocity_set.d == 3) else 1)
            table, n_bc = self.build_bc_table(tuple(f_0.shape[1:]))
            x_begin, x_end = (halo, dom.nx - halo) if self.velocity_set.d == 3 else (0, dom.nx)
            self._plan = (key, dom, table, n_bc, x_begin, x_end)
        return self._plan

    _LISTED_KINDS = (_lib.BC_DONOTHING, _lib.BC_ZOUHE_VELOCITY, _lib.BC_ZOUHE_PRESSURE, _lib.BC_EXTRAPOLATION,
                     _lib.BC_REGULARIZED_VELOCITY, _lib.BC_REGULARIZED_PRESSURE)

    CELL_LIST_MIN = 100_000

    def _bc_cell_list(self, f_0, bc_mask):
        """Sorted device list of the cells whose BC kind needs the per-cell engine (Zou-He, regularized, extrapolation
        outflow, do-nothing) plus its per-plane offsets, cached per ``bc_mask`` content.  ``xlb_step_ex`` then updates
        those cells with one dense launch over the list instead of diverging inside the fused kernel.  The reference
        specialises its kernel per BC list at construction (nse_stepper.py:351-380); this is the analogous setup step."""
        key = (bc_mask.data_ptr(), bc_mask._version, getattr(bc_mask, "_xlb_generation", None), tuple(bc_mask.shape))
        if self._cells is not None and self._cells[0] == key:
            return self._cells
        self._cells = (key, None, None)
        listed = any(int(bc.abi_kind) in self._LISTED_KINDS for bc in self.boundary_conditions)
        if not (self._use_cell_list and listed) or bc_mask.data_ptr() % 16 != 0:
            return self._cells
        _, dom, table, n_bc, _, _ = self._get_plan(f_0)
        count = torch.zeros(1, dtype=torch.int64, device=bc_mask.device)
        args = (byref(dom), bc_mask.data_ptr(), table.data_ptr(), 0, 0, int(dom.nx))
        check(lib().xlb_list_bc_cells(*args, None, 0, count.data_ptr(), stream_ptr()), "IncompressibleNavierStokesStepper")
        n = int(count.item())
        if n == 0 or (self._use_cell_list == "auto" and n < self.CELL_LIST_MIN):
            return self._cells
        cells = torch.empty(n, dtype=torch.int64, device=bc_mask.device)
        count.zero_()
        check(lib().xlb_list_bc_cells(*args, cells.data_ptr(), n, count.data_ptr(), stream_ptr()), "IncompressibleNavierStokesStepper")
        cells = torch.sort(cells).values
        plane = int(dom.ny) * int(dom.nz)
        edges = torch.arange(int(dom.nx) + 1, dtype=torch.int64, device=bc_mask.device) * plane
        offsets = torch.searchsorted(cells, edges).cpu().tolist()
        self._cells = (key, cells, offsets)
        return self._cells

    _NEEDS_MISSING = (_lib.BC_HALFWAY, _lib.BC_ZOUHE_VELOCITY, _lib.BC_ZOUHE_PRESSURE, _lib.BC_EXTRAPOLATION,
                      _lib.BC_REGULARIZED_VELOCITY, _lib.BC_REGULARIZED_PRESSURE)

    def _missing_bits(self, f_0, missing_mask):
        """``missing_mask`` packed to one uint32 per cell (``XlbCollisionParams.missing_bits``, ``xlb_pack_missing``) for the
        boundary paths that need a cell's missing directions: one load instead of q byte loads from q planes.  Built once per
        mask content (same cache rule as the BC cell list); ``backend_config["packed_missing"] = False`` turns it off."""
        key = (missing_mask.data_ptr(), missing_mask._version, getattr(missing_mask, "_xlb_generation", None), tuple(missing_mask.shape))
        if self._mbits is not None and self._mbits[0] == key:
            return self._mbits[1]
        self._mbits = (key, None)
        wanted = self.backend_config.get("packed_missing", True) if self.backend_config else True
        if os.environ.get("XLB_B200_PACKED_MISSING") in ("0", "1") and not (self.backend_config and "packed_missing" in self.backend_config):
            wanted = os.environ["XLB_B200_PACKED_MISSING"] == "1"  # experiments, like XLB_B200_BC_CELL_LIST
        if not wanted or not any(int(bc.abi_kind) in self._NEEDS_MISSING for bc in self.boundary_conditions):
            return None
        _, dom, _, _, _, _ = self._get_plan(f_0)
        bits = torch.empty(tuple(missing_mask.shape[1:]), dtype=torch.int32, device=missing_mask.device)
        check(lib().xlb_pack_missing(byref(dom), missing_mask.data_ptr(), bits.data_ptr(), stream_ptr()), "IncompressibleNavierStokesStepper")
        self._mbits = (key, bits)
        return bits

    def invalidate_bc_cells(self):
        """Forget the cached BC cell list and packed missing bits.  Needed only after ``bc_mask`` was rewritten behind torch's back (a raw-pointer
        write from foreign code); torch in-place ops and this package's maskers are noticed automatically."""
        self._cells = None
        self._mbits = None

    # ------------------------------------------------------------------ the step
    def launch_range(self, f_0, f_1, bc_mask, missing_mask, omega, timestep, x_begin, x_end):
        """One ``xlb_step`` launch that updates planes ``[x_begin, x_end)`` of ``f_1`` (local plane indices, ghost
        planes included in the numbering)."""
        rng = _lib.XlbStepRange(int(x_begin), 1, int(x_end) - int(x_begin), 0, 1, 0)
        self.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep, rng)

    def launch(self, f_0, f_1, bc_mask, missing_mask, omega, timestep, rng=None, halo=None):
        """One ``xlb_step_range`` launch: the planes / rows of ``rng`` (an ``XlbStepRange``; ``None`` = the stepper's whole
        block), optionally with the halo push ``halo`` (an ``XlbHaloPush``) of the outgoing edge populations into the ring
        neighbours' buffers.  The multi-GPU schedule (xlb_b200/distribute) issues the two edge planes of a slab as ONE
        strided launch with the push, and the interior as a second launch."""
        q = self.velocity_set.q
        if f_0.dtype != self.store_dtype or f_1.dtype != self.store_dtype:
            raise ValueError(f"populations must be stored as {self.store_dtype} under policy {self.precision_policy.name}")
        self._check_field(f_0, "f_0", q)
        self._check_field(f_1, "f_1", q)
        if f_0.shape != f_1.shape:
            raise ValueError("f_0 and f_1 must have the same shape")
        _, dom, table, n_bc, x_begin, x_end = self._get_plan(f_0)
        if rng is None:
            rng = self._whole_range(dom, x_begin, x_end)
        if n_bc:
            self._check_field(bc_mask, "bc_mask", 1)
            self._check_field(missing_mask, "missing_mask", q)
        cp = self._collision_params
        cp.bc_cells, cp.n_bc_cells = None, 0
        cp.missing_bits = None
        if n_bc:
            bits = self._missing_bits(f_0, missing_mask)
            if bits is not None:
                cp.missing_bits = bits.data_ptr()
        if n_bc and rng.n_planes > 0:
            _, cells, offsets = self._bc_cell_list(f_0, bc_mask)
            if cells is not None:
                # the list slice that covers the launch's planes; the kernel skips entries of planes / rows in between
                x_last = rng.x_begin + (rng.n_planes - 1) * rng.x_stride
                lo, hi = offsets[int(rng.x_begin)], offsets[int(x_last) + 1]
                if hi > lo:
                    cp.bc_cells, cp.n_bc_cells = cells.data_ptr() + 8 * lo, hi - lo
        rc = lib().xlb_step_range(
            byref(dom), self.collision.abi_model, byref(cp), byref(rng), byref(halo) if halo is not None else None, f_0.data_ptr(), f_1.data_ptr(),
            bc_mask.data_ptr() if n_bc else None, missing_mask.data_ptr() if n_bc else None, table.data_ptr() if n_bc else None, n_bc,
            float(omega), int(timestep), self.kernel_variant, stream_ptr(),
        )
        check(rc, "IncompressibleNavierStokesStepper")

    def _whole_range(self, dom, x_begin, x_end):
        """The owned cells of this block: all of it, or -- on a slab grid -- everything but the ghost planes (3-D) /
        ghost rows (D2Q9, whose fields run as nx = 1 with the user's x axis on the kernel's y axis)."""
        halo = getattr(self.grid, "halo", 0)
        if self.velocity_set.d == 2 and halo:
            return _lib.XlbStepRange(0, 1, 1, halo, 1, int(dom.ny) - 2 * halo)
        return _lib.XlbStepRange(int(x_begin), 1, int(x_end) - int(x_begin), 0, 1, 0)

    def launch_in_place(self, f, bc_mask, omega):
        """One ``xlb_step_inplace`` launch on the single buffer ``f``; alternates P and L steps (``aa_parity``)."""
        q = self.velocity_set.q
        if f.dtype != self.store_dtype:
            raise ValueError(f"populations must be stored as {self.store_dtype} under policy {self.precision_policy.name}")
        self._check_field(f, "f", q)
        _, dom, table, n_bc, _, _ = self._get_plan(f)
        if n_bc:
            self._check_field(bc_mask, "bc_mask", 1)
        rc = lib().xlb_step_inplace(
            byref(dom), self.collision.abi_model, byref(self._collision_params), f.data_ptr(), bc_mask.data_ptr() if n_bc else None,
            table.data_ptr() if n_bc else None, n_bc, float(omega), int(self.aa_parity), stream_ptr(),
        )
        check(rc, "IncompressibleNavierStokesStepper")
        self.aa_parity ^= 1

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f_0, f_1, bc_mask, missing_mask, omega, timestep):
        if self.in_place:
            if f_0.data_ptr() != f_1.data_ptr():
                raise ValueError("this stepper was built with in_place=True: pass the one buffer prepare_fields() returned as f_0 and f_1")
            self.launch_in_place(f_0, bc_mask, omega)
            return f_0, f_1
        self.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep)
        return f_0, f_1
