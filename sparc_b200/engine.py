"""Batched engine: owns the device state blob (torch memory) and drives the C-ABI.

PyTorch is plumbing here — device allocation, streams, pinned host buffers.  All arithmetic
happens in the sm_100a kernels behind include/sparc_b200.h.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np
import torch

from . import _native as N
from .params import SparcParams

_TORCH_DTYPE = {"d": torch.float64, "f": torch.float32, "i": torch.int32}
_ELEM = {"d": 8, "f": 4, "i": 4}


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class Engine:
    """State of `n_envs` environments on one CUDA device + the step/observe/reset calls."""

    def __init__(self, n_envs: int, params: SparcParams, device="cuda:0", seed: int = 0,
                 env_id_base: int = 0):
        if not torch.cuda.is_available():
            raise N.SparcError("sparc_b200 needs a CUDA device (there is no CPU fallback)")
        self.lib = N.lib()
        self.n_envs = int(n_envs)
        self.params = params
        self.n_segments = int(params.n_segments)
        self.device = torch.device(device)
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.env_id_base = int(env_id_base)
        self._ctrl = None
        nbytes = self.lib.sparc_state_bytes(self.n_envs, self.n_segments)
        if nbytes == 0:
            raise N.SparcError("bad n_envs / n_segments")
        with torch.cuda.device(self.device):
            self.blob = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
            self.n_pad = self.lib.sparc_padded_segments(self.n_segments)
            self.fields: Dict[str, torch.Tensor] = {}
            for name, dt, off, cnt in N.field_table(self.n_envs, self.n_segments):
                view = self.blob[off: off + cnt * _ELEM[dt]].view(_TORCH_DTYPE[dt])
                if name == "T":
                    view = view.view(self.n_envs, self.n_pad)
                self.fields[name] = view
            # action staging (device) + pinned host mirrors for the host-buffer path
            self._act_dev = {
                "servo": torch.zeros(self.n_envs, dtype=torch.float32, device=self.device),
                "target_voltage": torch.zeros(self.n_envs, dtype=torch.float32, device=self.device),
                "current_mode": torch.zeros(self.n_envs, dtype=torch.int32, device=self.device),
                "on_time": torch.zeros(self.n_envs, dtype=torch.float32, device=self.device),
                "off_time": torch.zeros(self.n_envs, dtype=torch.float32, device=self.device),
            }
            self.obs_dev = torch.zeros(N.N_OBS, self.n_envs, dtype=torch.float32, device=self.device)
            self.stats_dev = torch.zeros(N.N_STATS, dtype=torch.float64, device=self.device)
            self._scratch = torch.zeros(self.lib.sparc_host_scratch_bytes(self.n_envs), dtype=torch.uint8,
                                        device=self.device)
            # one pinned buffer, five arrays at the library's pitch: sparc_step_host moves them in one copy
            pitch = int(self.lib.sparc_host_action_pitch(self.n_envs))
            self._act_host_buf = torch.zeros(5 * pitch, dtype=torch.uint8).pin_memory()
            self._act_host = {k: self._act_host_buf[i * pitch: i * pitch + 4 * self.n_envs].view(v.dtype)
                              for i, (k, v) in enumerate(self._act_dev.items())}
            self.obs_host = torch.zeros(N.N_OBS, self.n_envs, dtype=torch.float32).pin_memory()
            N.check(self.lib.sparc_init(self.blob.data_ptr(), C.byref(self.params), self.n_envs, self._stream()))

    # ------------------------------------------------------------------------------------------
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def __getitem__(self, name: str) -> torch.Tensor:
        return self.fields[name]

    @property
    def T(self) -> torch.Tensor:
        """Wire temperature field, [n_envs, n_segments] float32 view of the live state."""
        return self.fields["T"][:, : self.n_segments]

    def launch_info(self) -> dict:
        out = (C.c_int32 * 4)()
        N.check(self.lib.sparc_launch_info(self.n_envs, self.n_segments, out))
        return dict(lanes_per_env=out[0], cells_per_lane=out[1], block_threads=out[2], blocks=out[3])

    def reset(self, mask: Optional[torch.Tensor] = None, reseed: bool = False, reference_semantics: bool = False) -> None:
        """Fresh-env state for the masked envs (all when mask is None).  Each reset starts a new
        episode (a new Philox stream); `reseed=True` — the caller passed a seed — starts the episode
        counter over, so that `reset(seed=s)` replays the same episode every time."""
        m = None
        if mask is not None:
            m = mask.to(device=self.device, dtype=torch.uint8).contiguous()
        # reference_semantics: only EDMState is re-created, module-internal state survives (sparc_reset_reference)
        fn = self.lib.sparc_reset_reference if reference_semantics else (self.lib.sparc_reseed if reseed else self.lib.sparc_reset)
        with torch.cuda.device(self.device):
            N.check(fn(self.blob.data_ptr(), C.byref(self.params), self.n_envs, _ptr(m), self._stream()))

    def set_actions(self, servo, target_voltage, current_mode, on_time, off_time) -> None:
        """Copy one batched action (any array-likes broadcastable to [n_envs]) to the device."""
        for key, val in (("servo", servo), ("target_voltage", target_voltage), ("current_mode", current_mode),
                         ("on_time", on_time), ("off_time", off_time)):
            dst = self._act_dev[key]
            if isinstance(val, torch.Tensor):
                src = val.reshape(-1) if val.numel() > 1 else val.reshape(())
                dst.copy_(src.to(dtype=dst.dtype), non_blocking=True)
            else:
                arr = np.asarray(val)
                if arr.size == 1:
                    dst.fill_(arr.reshape(()).item())
                else:
                    dst.copy_(torch.from_numpy(np.ascontiguousarray(arr.reshape(-1))).to(dst.dtype), non_blocking=True)

    def set_controller(self, kind: int = 0, desired_gap: float = 0.0, gain: float = 0.0, limit: float = -1.0) -> None:
        """In-kernel controller for `step()` (include/sparc_b200.h SparcController): kind 1 = proportional
        gap controller evaluated at the start of every call instead of reading the staged servo array;
        kind 0 switches it off.  limit < 0 (or None): no clipping."""
        self._ctrl = None if not kind else N.SparcController(int(kind), 0, float(desired_gap), float(gain),
                                                             -1.0 if limit is None else float(limit))

    def _actions_struct(self, bufs) -> N.SparcActions:
        return N.SparcActions(bufs["servo"].data_ptr(), bufs["target_voltage"].data_ptr(),
                              bufs["current_mode"].data_ptr(), bufs["on_time"].data_ptr(),
                              bufs["off_time"].data_ptr())

    def step(self, n_us: int, draws: Optional[torch.Tensor] = None, observe: bool = False):
        """Advance every env by n_us reference step() calls with the staged device actions.

        draws: optional [n_envs, n_us, 5] float64 device tensor of injected random draws.
        observe=True also returns the [8, n_envs] observation tensor of `observe()`, written by the
        step kernel itself (no second launch)."""
        a = self._actions_struct(self._act_dev)
        dptr = None
        if draws is not None:
            assert draws.dtype == torch.float64 and draws.is_contiguous() and draws.device == self.blob.device
            assert tuple(draws.shape) == (self.n_envs, int(n_us), N.N_DRAW_SLOTS)
            dptr = draws.data_ptr()
        with torch.cuda.device(self.device):
            if getattr(self, "_ctrl", None) is not None:
                N.check(self.lib.sparc_step_ex(self.blob.data_ptr(), C.byref(self.params), C.byref(a), C.byref(self._ctrl),
                                               self.n_envs, int(n_us), self.seed, self.env_id_base, dptr,
                                               self.obs_dev.data_ptr() if observe else None, self._stream()))
                return self.obs_dev if observe else None
            if observe:
                N.check(self.lib.sparc_step_observe(self.blob.data_ptr(), C.byref(self.params), C.byref(a), self.n_envs,
                                                    int(n_us), self.seed, self.env_id_base, dptr,
                                                    self.obs_dev.data_ptr(), self._stream()))
                return self.obs_dev
            if draws is None:
                N.check(self.lib.sparc_step(self.blob.data_ptr(), C.byref(self.params), C.byref(a), self.n_envs,
                                            int(n_us), self.seed, self.env_id_base, self._stream()))
            else:
                N.check(self.lib.sparc_step_injected(self.blob.data_ptr(), C.byref(self.params), C.byref(a),
                                                     self.n_envs, int(n_us), dptr, self._stream()))
        return None

    def step_logged(self, n_us: int, log_every: int, n_log_envs: Optional[int] = None,
                    draws: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """step() that also samples the wire temperature field of envs [0, n_log_envs) every
        `log_every` µs: returns [n_log_envs, n_us // log_every, n_segments] float32 (device), the
        reference logger's (n_timesteps, n_segments) layout per env (utils/logger.py:128-150)."""
        n_log = self.n_envs if n_log_envs is None else int(n_log_envs)
        rows = int(n_us) // int(log_every)
        a = self._actions_struct(self._act_dev)
        with torch.cuda.device(self.device):
            if out is None:
                out = torch.zeros(n_log, rows, self.n_segments, dtype=torch.float32, device=self.device)
            assert out.is_contiguous() and tuple(out.shape) == (n_log, rows, self.n_segments)
            dptr = None
            if draws is not None:
                assert draws.dtype == torch.float64 and draws.is_contiguous() and draws.device == self.blob.device
                assert tuple(draws.shape) == (self.n_envs, int(n_us), N.N_DRAW_SLOTS)
                dptr = draws.data_ptr()
            N.check(self.lib.sparc_step_logged(self.blob.data_ptr(), C.byref(self.params), C.byref(a), self.n_envs,
                                               int(n_us), self.seed, self.env_id_base, dptr, out.data_ptr(),
                                               int(log_every), n_log, self._stream()))
        return out

    def step_sampled(self, n_us: int, sample_every: int, draws: Optional[torch.Tensor] = None,
                     log_every: int = 0, n_log_envs: Optional[int] = None):
        """step() with the device-side signal logger: returns (signals, tlog) where signals is
        [n_us // sample_every, N_SIGNALS, n_envs] float64 (rows of `_native.SIGNALS`, the state after
        every `sample_every`-th µs, like SimulationLogger.collect after that step(), utils/logger.py:118-160)
        and tlog the temperature log of step_logged (None unless log_every > 0)."""
        every = int(sample_every)
        assert every > 0
        rows = int(n_us) // every
        a = self._actions_struct(self._act_dev)
        with torch.cuda.device(self.device):
            sig = torch.zeros(rows, N.N_SIGNALS, self.n_envs, dtype=torch.float64, device=self.device)
            tlog, n_log = None, 0
            if log_every:
                n_log = self.n_envs if n_log_envs is None else int(n_log_envs)
                tlog = torch.zeros(n_log, int(n_us) // int(log_every), self.n_segments, dtype=torch.float32,
                                   device=self.device)
            dptr = None
            if draws is not None:
                assert draws.dtype == torch.float64 and draws.is_contiguous() and draws.device == self.blob.device
                assert tuple(draws.shape) == (self.n_envs, int(n_us), N.N_DRAW_SLOTS)
                dptr = draws.data_ptr()
            N.check(self.lib.sparc_step_sampled(self.blob.data_ptr(), C.byref(self.params), C.byref(a), self.n_envs,
                                                int(n_us), self.seed, self.env_id_base, dptr,
                                                tlog.data_ptr() if tlog is not None else None, int(log_every), n_log,
                                                sig.data_ptr(), every, self._stream()))
        return sig, tlog

    def step_host(self, n_us: int) -> torch.Tensor:
        """Host-buffer path: pinned host actions -> device, n_us steps, observation -> pinned host.

        Fill `engine.host_actions[...]` first.  Synchronises the stream; returns obs_host [8, n_envs]."""
        a = self._actions_struct(self._act_host)
        with torch.cuda.device(self.device):
            N.check(self.lib.sparc_step_host(self.blob.data_ptr(), C.byref(self.params), C.byref(a), self.n_envs,
                                             int(n_us), self.seed, self.env_id_base, self._scratch.data_ptr(),
                                             self.obs_host.data_ptr(), self._stream()))
        return self.obs_host

    @property
    def host_actions(self) -> Dict[str, torch.Tensor]:
        return self._act_host

    def host_stepper(self, n_us: int):
        """`step_host(n_us)` with every ctypes argument built once: returns a zero-argument callable that
        runs one `sparc_step_host` (H2D of the pinned action arrays, n_us steps, D2H of the observation,
        stream sync) and leaves the result in `obs_host`.  For per-µs callers (the drop-in
        `WireEDMEnv.step`) the per-call Python overhead of `step_host` is larger than the GPU work.
        The stream and seed are captured now; the caller's current device must be this engine's."""
        a = self._actions_struct(self._act_host)
        fn = self.lib.sparc_step_host
        if torch.cuda.current_device() != (self.device.index or 0):
            return lambda: self.step_host(n_us)
        args = (C.c_void_p(self.blob.data_ptr()), C.byref(self.params), C.byref(a), self.n_envs, int(n_us),
                C.c_uint64(self.seed), C.c_uint32(self.env_id_base), C.c_void_p(self._scratch.data_ptr()),
                C.c_void_p(self.obs_host.data_ptr()), C.c_void_p(self._stream()))

        def call(_fn=fn, _args=args, _keep=a):
            rc = _fn(*_args)
            if rc:
                N.check(rc)
        return call

    def observe(self) -> torch.Tensor:
        """[8, n_envs] float32 device tensor: gap, wire velocity, mean voltage of the last call,
        sparks and short-circuit µs in the last call, max wire temperature, debris density, flags."""
        with torch.cuda.device(self.device):
            N.check(self.lib.sparc_observe(self.blob.data_ptr(), self.n_envs, self.n_segments,
                                           self.obs_dev.data_ptr(), self._stream()))
        return self.obs_dev

    def reduce_stats(self) -> torch.Tensor:
        """fp64[8] device tensor: sparks, short µs, short events, crater µm³, short pulses,
        env-µs simulated, broken envs, target-reached envs (summed over this engine's envs)."""
        with torch.cuda.device(self.device):
            N.check(self.lib.sparc_reduce_stats(self.blob.data_ptr(), self.n_envs, self.n_segments,
                                                self.stats_dev.data_ptr(), self._stream()))
        return self.stats_dev

    def bytes_per_control_step(self) -> int:
        """Algorithmic HBM bytes one step call moves: the whole state read once and written once."""
        scratch = sum(int(self.fields[k].numel()) * self.fields[k].element_size()
                      for k in ("mech_traj", "mech_ckpt", "mech_prog") if k in self.fields)
        return 2 * (int(self.blob.numel()) - scratch)


def measure_fp32_peak(device="cuda:0") -> float:
    """Measured FP32 FMA peak [TFLOP/s] of `device` (roofline denominator for the step kernel)."""
    out = C.c_double(0.0)
    with torch.cuda.device(device):
        scratch = torch.empty(8 << 20, dtype=torch.uint8, device=device)
        N.check(N.lib().sparc_measure_fp32_peak(scratch.data_ptr(), scratch.numel(), C.byref(out)))
    return float(out.value)
