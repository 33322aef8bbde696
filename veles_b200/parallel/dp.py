"""Synchronous data parallelism over the GPUs of one NVSwitch box.

Every rank (one process per GPU) runs the identical workflow on its shard of each global
minibatch.  What the reference did with a ZeroMQ parameter-server star — pickled weights down,
pickled gradients up, merged by one Python master under a lock
(``veles/server.py:141-169, 401-430``, ``veles/client.py:138-160``, ``veles/workflow.py:456-573``) —
is ONE device kernel per bucket of layers here (``csrc/comm_fused.cu``): gradient all-reduce through NVLS
multicast (or P2P loads), the Znicz GD update in registers, and the updated fp32 weights + bf16
shadow multicast-stored back to every rank.  Layers are bucketed in backward order (a bucket closes once it
holds ``bucket_bytes`` of parameters: AlexNet = fc8 | fc7 | fc6 | conv5+conv4 | conv3+conv2 | conv1); a bucket's
kernel is launched on a side stream as soon as its last layer's backward GEMMs are enqueued and runs on a few SMs
that the backward GEMMs leave free for it (``vb_set_sm_reserve``), so the exchange overlaps the rest of the
backward pass instead of taking turns with it.

``mode="nccl"`` is the conventional baseline (NCCL all-reduce, then the local GD kernel).
``mode="host"`` runs the same protocol for the numpy backend (gradients all-reduced as host arrays over whatever
backend the process group has, normally gloo): the CPU oracle of the fused kernel's math and the way the host-side
logic (loader sharding, counter reductions, scheduling) is tested without GPUs.

torch.distributed is used only for plumbing: rendezvous of symmetric-memory slabs
(``torch.distributed._symmetric_memory``), epoch-boundary counter sums, barriers.
"""
from __future__ import annotations

import ctypes
from ctypes import c_double, c_float, c_int, c_longlong, c_void_p

import numpy

SYMMETRIC_KINDS = ("weights", "weights_bf16", "grads", "grads1", "grads_in", "dp_flags", "dp_counters", "velocity")
COUNTER_SLOT = 16384   # doubles per parity of the device-side counter exchange (csrc/comm_counters.cu)
MAX_WORLD = 8
MAX_SEGS = 16          # csrc/comm_fused.cu
SEG_ZERO_OTHER = 1     # the segment's gradient kernels accumulate: the update kernel clears the other parity's buffer
SEG_BCAST_MASTER = 2   # forward passes read this segment's fp32 master: always publish it to every rank


class DataParallel(object):
    def __init__(self, mode="fused", n_ctas=None, timeout_s=20.0, use_multimem=None, shard_master=True,
                 weight_delivery="multicast", split_handshake=False, bucket_bytes=4 << 20, exchange="pull"):
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("DataParallel needs an initialised torch.distributed process group")
        self.mode = mode
        self.rank = dist.get_rank()
        self.world_size = dist.get_world_size()
        if self.world_size > MAX_WORLD:
            raise ValueError("the fused kernel supports at most %d ranks" % MAX_WORLD)
        # CTAs of the update kernel = SMs the backward GEMMs leave free while an exchange is pending.  One SM pulls
        # ~8 GB/s through multimem.ld_reduce (or peer loads: the same), whatever it keeps in flight, and each rank
        # reduces 244 MB / world for AlexNet.  Measured at 2 ranks (ms per step, 1 GPU = 2.32): 8 SMs 3.45, 16 SMs
        # 2.49, 24 SMs 2.57, 32 SMs 2.56 -> 16 SMs up to 4 ranks, 8 beyond.
        self.n_ctas = int(n_ctas) if n_ctas else (16 if self.world_size <= 4 else 8)
        self.n_ctas = max(2, (self.n_ctas + 1) & ~1)
        # peer-TMA delivery describes ONE layer's weight matrix per launch: no bucketing there
        self.bucket_bytes = 0 if weight_delivery == "peer_tma" else int(bucket_bytes)
        # exchange: how gradients reach the rank that owns (updates) a slice:
        #   "pull": the owner reads them: multimem.ld_reduce through the switch (or peer loads), the next batch
        #           requested while the current one is consumed (default);
        #   "push": every rank STORES its contributions into the owner's staging slab, the owner reduces from local
        #           memory.  Measured at 2 ranks with 8 SMs: slower than the pull form (fc6: 1.9 vs 1.2 ms) - remote
        #           stores from a few SMs do not reach link bandwidth either; kept as an option.
        if exchange not in ("push", "pull"):
            raise ValueError("exchange must be 'push' or 'pull'")
        self.exchange = exchange
        self.profile = False       # eager runs: time every exchange kernel with CUDA events (bucket_times())
        self._prof_events = []
        self.buckets = []          # launch units: lists of GD units exchanged by one kernel, in backward order
        self.bucket_of = {}
        self._reserved = False
        self.auto_finish = False      # set by the launcher for graph-run workflows (FlatStep calls finish_step itself)
        self.last_unit = None         # the exchanged GD unit that runs last in a backward pass
        self.timeout_s = timeout_s
        self.use_multimem = use_multimem
        # shard_master: each rank keeps only ITS slice of the fp32 master weights current (the slice it updates);
        # the full master is reassembled by sync_master() at class/epoch boundaries, when the host may look.
        self.shard_master = bool(shard_master)
        # weight_delivery: how the updated bf16 operand weights reach the other ranks:
        #   "multicast": multimem.st of every updated slice into all replicas (inside the update kernel);
        #   "peer_tma":  nothing is sent; forward / err_input GEMMs TMA-load each weight tile from its owner.
        if weight_delivery not in ("multicast", "peer_tma"):
            raise ValueError("weight_delivery must be 'multicast' or 'peer_tma'")
        self.weight_delivery = weight_delivery
        # split_handshake: a 1-CTA launch waits for the peers' "gradients ready" flags, THEN the worker CTAs start,
        # so rank skew never parks 64 CTAs on SMs the backward GEMMs could use
        self.split_handshake = bool(split_handshake)
        self.master_stale = False
        self.parity = 0
        self.grad_scale = 1.0      # 1/world when every rank normalised by its LOCAL batch size
        self.units = []            # GD units in registration order -> channel index
        self.unit_info = {}
        self.handles = {}          # id(slab tensor) -> symm handle
        self.symm_slabs = []       # [(kind, tensor)] in allocation order (identical on every rank)
        self.comm_stream = None
        self.arena = None
        self.wf = None
        self._dist = dist
        self._pending_events = []

    # -- setup -----------------------------------------------------------------------------------------------
    def __reduce__(self):
        # process groups, symmetric-memory handles and streams belong to this process: a snapshot of a workflow that
        # references its group (Decision.dp_group) stores None there; the launcher attaches a new group on resume
        return (type(None), ())

    def attach(self, wf):
        """Before ``wf.initialize``: route the exchanged arena kinds into symmetric memory."""
        self.wf = wf
        if self.mode not in ("fused", "nccl", "host"):
            raise ValueError("DataParallel mode must be 'fused', 'nccl' or 'host'")
        if self.mode == "fused":
            wf.arena_allocator_ = self._allocate
        wf.dp_group_ = self

    def _allocate(self, kind, nbytes):
        if kind not in SYMMETRIC_KINDS:
            return None
        import torch
        import torch.distributed._symmetric_memory as symm_mem
        t = symm_mem.empty(nbytes, dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
        self.symm_slabs.append((kind, t))
        return t

    def finalize(self, wf):
        """After ``wf.initialize``: exchange handles, describe every GD unit's segments."""
        import torch
        import torch.distributed._symmetric_memory as symm_mem
        dev = wf.device
        if self.mode == "host":
            for gd in wf.gds:
                if hasattr(gd, "gradient_weights") and gd.weights is not None and gd.weights:
                    gd.dp_group_ = self
                    self.units.append(gd)
            self.last_unit = self.units[0] if self.units else None      # wf.gds is in forward order
            if hasattr(wf, "decision") and wf.decision is not None:
                wf.decision.dp_group = self
            return
        # high priority: when SMs free up, the exchange kernel's CTAs are placed before the (many) pending CTAs of the
        # backward kernels - otherwise it can starve behind them until the backward pass ends
        self.comm_stream = torch.cuda.Stream(device=dev.torch_device, priority=-1)
        self.arena = wf.get_arena(dev)
        for gd in wf.gds:
            if hasattr(gd, "gradient_weights") and gd.weights is not None and gd.weights:
                gd.dp_group_ = self
                self.units.append(gd)
        self.last_unit = self.units[0] if self.units else None          # wf.gds is in forward order
        if hasattr(wf, "decision") and wf.decision is not None:
            wf.decision.dp_group = self
        if self.mode != "fused":
            return
        if len(self.units) > 64:
            raise ValueError("at most 64 trainable layers are supported by the flag layout")
        arena = self.arena
        # second gradient parity + flag words, mirrored/allocated identically on every rank
        for gd in self.units:
            for w in (gd.weights, gd.bias):
                if w is not None and w:
                    # pull: second gradient parity (peers read my slab while I start the next step);
                    # push: staging slab the peers store their contributions into
                    arena.mirror("grads1" if self.exchange == "pull" else "grads_in", w.devmem.view(-1))
        self.flags = arena.alloc("dp_flags", "flags", 64 * 2 * MAX_WORLD, torch.int32)
        # class / epoch ends: counters are summed and the sharded state gathered by kernels over peer memory
        from ..ops import _lib
        L = _lib.lib()
        L.vb_counter_region_bytes.restype = ctypes.c_longlong
        self.counter_region = arena.alloc("dp_counters", "counters",
                                          int(L.vb_counter_region_bytes(ctypes.c_longlong(COUNTER_SLOT))), torch.uint8)
        self.counter_epoch = torch.zeros(1, dtype=torch.int32, device=dev.torch_device)
        self.counter_in = torch.zeros(COUNTER_SLOT, dtype=torch.float64, device=dev.torch_device)
        self.counter_out = torch.zeros(COUNTER_SLOT, dtype=torch.float64, device=dev.torch_device)
        self.counter_host = torch.zeros(COUNTER_SLOT, dtype=torch.float64).pin_memory()
        self.epochs = torch.zeros(64, dtype=torch.int32, device=dev.torch_device)
        self.cta_counter = torch.zeros(128, dtype=torch.int32, device=dev.torch_device)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev.torch_device)
        torch.cuda.synchronize()
        group = self._dist.group.WORLD
        for kind, t in self.symm_slabs:
            hdl = symm_mem.rendezvous(t, group=group)
            self.handles[t.data_ptr()] = hdl
        if self.use_multimem is None:
            self.use_multimem = all(int(h.multicast_ptr) != 0 for h in self.handles.values())
        for gd in self.units:
            self.unit_info[id(gd)] = self._describe(gd)
        self._build_buckets()
        cidx, coff = arena.locate(self.counter_region, "dp_counters")
        cl = self._slab_ptrs("dp_counters", cidx)
        self.counter_ptrs = (cl[0] + coff, [p + coff for p in cl[2]])
        self._dist.barrier()
        torch.cuda.synchronize()

    def _build_buckets(self):
        """Group the exchanged units, in backward order, into launch buckets (one flag channel each)."""
        self.buckets, self.bucket_of = [], {}
        cur = None
        for gd in reversed(self.units):          # wf.gds is in forward order: the backward pass meets them reversed
            info = self.unit_info[id(gd)]
            if (cur is None or cur["bytes"] >= self.bucket_bytes or cur["slab"] != info["slab"]
                    or cur["shadow_slab"] != info["shadow_slab"] or len(cur["segs"]) + len(info["segs"]) > MAX_SEGS):
                cur = {"channel": len(self.buckets), "units": [], "segs": [], "kinds": [], "bytes": 0,
                       "slab": info["slab"], "shadow_slab": info["shadow_slab"], "pending": 0}
                for key in ("grads", "weights", "shadow", "vel", "flags", "stage"):
                    cur[key] = info[key]
                self.buckets.append(cur)
            cur["units"].append(gd)
            cur["segs"] += info["segs"]
            cur["kinds"] += [(gd, k) for k in info["kinds"]]
            cur["bytes"] += 4 * sum(int(sg[1]) for sg in info["segs"])
            cur["peer_shadow"] = cur.get("peer_shadow", False) or bool(info.get("peer_shadow"))
            self.bucket_of[id(gd)] = cur
        if len(self.buckets) > 64:
            raise ValueError("at most 64 exchange buckets are supported by the flag layout")
        for b in self.buckets:
            # small buckets are latency-bound: two CTAs do; the reserve is sized for the largest launch
            b["n_ctas"] = self.n_ctas if b["bytes"] >= (1 << 20) else 2
            if b["peer_shadow"] and len(b["units"]) > 1:
                raise ValueError("peer-TMA weight delivery needs one layer per bucket (bucket_bytes=0)")

    def _slab_ptrs(self, kind, idx):
        """(local base, multicast base or 0, [peer bases]) of slab ``idx`` of ``kind``."""
        t = self.arena.slabs[kind][idx][0]
        hdl = self.handles[t.data_ptr()]
        delta = t.data_ptr() - int(hdl.buffer_ptrs[self.rank])
        peers = [int(p) + delta for p in hdl.buffer_ptrs]
        mc = int(hdl.multicast_ptr)
        return t.data_ptr(), (mc + delta) if mc else 0, peers

    def _describe(self, gd):
        import torch
        arena = self.arena
        fwd = gd.forward
        segs, hyper_keys = [], []
        slab_idx = None
        peer_shadow = None
        for w, is_bias in ((gd.weights, False), (gd.bias, True)):
            if w is None or not w:
                continue
            idx, off = arena.locate(w.devmem)
            if slab_idx is None:
                slab_idx = idx
            elif idx != slab_idx:
                raise ValueError("weights and bias of %s live in different arena slabs" % gd.name)
            shadow_off, row_len, row_pitch = -1, 1, 1
            if not is_bias and fwd is not None and fwd.weights_bf16_ is not None:
                sh = fwd.weights_bf16_
                sidx, soff = arena.locate(sh, "weights_bf16")
                if sidx != 0 and len(arena.slabs["weights_bf16"]) > 1:
                    pass
                self._shadow_slab = sidx
                shadow_off, row_len, row_pitch = soff // 2, sh.shape[1], sh.stride(0)
            flags = 0
            if is_bias or not getattr(gd, "dw_store_", False):
                flags |= SEG_ZERO_OTHER            # column sums / split-K weight gradients accumulate into the buffer
            if shadow_off < 0 or (hasattr(gd, "tf32_mode") and gd.tf32_mode()):
                flags |= SEG_BCAST_MASTER          # the forward pass reads the fp32 master of this segment
            segs.append([off // 4, w.size, shadow_off, row_len, row_pitch, flags])
            hyper_keys.append("bias" if is_bias else "weights")
            if (not is_bias and shadow_off >= 0 and self.weight_delivery == "peer_tma" and row_len == row_pitch
                    and getattr(fwd, "PEER_WEIGHTS_OK", False) and not getattr(fwd, "weights_transposed", False)):
                n_rows = fwd.weights_bf16_.shape[0]
                W = self.world_size
                if n_rows % W == 0 and (n_rows // W) % 64 == 0 and (n_rows // W * row_len) % 4 == 0:
                    peer_shadow = (sidx, soff, n_rows // W)
        info = {"slab": slab_idx, "segs": segs, "kinds": hyper_keys,
                "shadow_slab": getattr(self, "_shadow_slab", 0)}
        if self.exchange == "push":
            info["grads"] = [self._slab_ptrs("grads", slab_idx)] * 2
            info["stage"] = self._slab_ptrs("grads_in", slab_idx)
        else:
            info["grads"] = [self._slab_ptrs("grads", slab_idx), self._slab_ptrs("grads1", slab_idx)]
            info["stage"] = (0, 0, [0] * self.world_size)
        info["weights"] = self._slab_ptrs("weights", slab_idx)
        info["shadow"] = self._slab_ptrs("weights_bf16", info["shadow_slab"])
        if peer_shadow is not None:
            # the update kernel keeps this layer's bf16 shard local; its GEMMs read every tile from the owner
            sidx, soff, shard_rows = peer_shadow
            _, _, peers = self._slab_ptrs("weights_bf16", sidx)
            fwd.peer_weights_ = ([p + soff for p in peers], shard_rows, self.rank)
            info["peer_shadow"] = True
        info["vel"] = arena.slabs["velocity"][slab_idx][0].data_ptr()
        info["vel_symmetric"] = arena.slabs["velocity"][slab_idx][0].data_ptr() in self.handles
        fl_idx, fl_off = arena.locate(self.flags, "dp_flags")
        fl = self._slab_ptrs("dp_flags", fl_idx)
        info["flags"] = (fl[0] + fl_off, [p + fl_off for p in fl[2]])
        # per-parity views for the backward GEMMs
        views = []
        for kind in (("grads", "grads") if self.exchange == "push" else ("grads", "grads1")):
            slab = arena.slabs[kind][slab_idx][0]
            pair = []
            for w in (gd.weights, gd.bias):
                if w is None or not w:
                    pair.append(None)
                    continue
                _, off = arena.locate(w.devmem)
                pair.append(slab[off:off + w.size * 4].view(torch.float32))
            views.append(tuple(pair))
        info["views"] = views
        return info

    # -- per step ---------------------------------------------------------------------------------------------------
    def grads_of(self, gd):
        if self.mode != "fused":
            return (gd.gradient_weights.devmem, gd.gradient_bias.devmem if (gd.bias is not None and gd.bias) else None)
        return self.unit_info[id(gd)]["views"][self.parity]

    def update_unit(self, gd):
        import torch
        if self.mode == "fused":
            res = self._update_fused(gd)
            self._maybe_finish(gd)
            return res
        # --- NCCL baseline: all-reduce, then the single-GPU fused GD kernel
        dist = self._dist
        op = dist.ReduceOp.AVG if self.grad_scale != 1.0 else dist.ReduceOp.SUM
        dist.all_reduce(gd.gradient_weights.devmem, op=op)
        if gd.bias is not None and gd.bias:
            dist.all_reduce(gd.gradient_bias.devmem, op=op)
        group, gd.dp_group_ = gd.dp_group_, None
        try:
            gd.device_update()
        finally:
            gd.dp_group_ = group
        self._maybe_finish(gd)

    def _maybe_finish(self, gd):
        """Workflows driven by their own unit graph (``wf.run()``, the CLI) have no stepper that calls
        ``finish_step``: the unit that runs last in the backward chain closes the step itself."""
        if self.auto_finish and gd is self.last_unit:
            self.finish_step()

    def _update_fused(self, gd):
        """The unit's gradients are enqueued on the current stream: note it, and launch its bucket's exchange kernel
        once the bucket's last unit (in backward order) has reported."""
        b = self.bucket_of[id(gd)]
        b["pending"] += 1
        self.master_stale = True      # at least the momentum now lives on its owners only
        gd.weights.mark_device_written()
        gd.gradient_weights_with_moment.mark_device_written()
        if gd.bias is not None and gd.bias:
            gd.bias.mark_device_written()
            gd.gradient_bias_with_moment.mark_device_written()
        if gd is b["units"][-1]:
            self._launch_bucket(b)

    def _launch_bucket(self, b):
        import torch
        from ..ops import _lib, kernels
        b["pending"] = 0
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self.comm_stream.wait_event(ev)
        if not self._reserved:
            # GEMMs enqueued from here to finish_step leave n_ctas SMs to the exchange kernels
            _lib.lib().vb_set_sm_reserve(c_int(self.n_ctas))
            self._reserved = True
        nseg = len(b["segs"])
        segs = (c_longlong * (6 * nseg))(*[int(v) for s in b["segs"] for v in s])
        hyper = []
        for gd, kind in b["kinds"]:
            if kind == "weights":
                lr, mom = gd.eff(gd.learning_rate, gd.gradient_moment) if hasattr(gd, "eff") else (
                    gd.learning_rate, gd.gradient_moment)
                hyper += [lr, gd.weights_decay, gd.l1_vs_l2, mom]
            else:
                lr, mom = gd.eff(gd.learning_rate_bias, gd.gradient_moment_bias) if hasattr(gd, "eff") else (
                    gd.learning_rate_bias, gd.gradient_moment_bias)
                hyper += [lr, gd.weights_decay_bias, gd.l1_vs_l2_bias, mom]
        hyper = (c_float * (4 * nseg))(*hyper)
        push = self.exchange == "push"
        g_local, g_mc, g_peers = b["grads"][self.parity]
        go_local = 0 if push else b["grads"][1 - self.parity][0]
        st_local, _, st_peers = b["stage"]
        w_local, w_mc, w_peers = b["weights"]
        s_local, s_mc, s_peers = b["shadow"]
        f_local, f_peers = b["flags"]
        W = self.world_size

        def arr(ptrs):
            return (c_void_p * W)(*[c_void_p(p) for p in ptrs])

        prof = self.profile and not torch.cuda.is_current_stream_capturing()
        with torch.cuda.stream(self.comm_stream):
            if prof:
                t0 = torch.cuda.Event(enable_timing=True)
                t0.record(self.comm_stream)
            rc = _lib.lib().vb_fused_allreduce_gd(
                c_int(W), c_int(self.rank), c_int(b["channel"]), c_int(nseg), segs, hyper,
                c_void_p(g_local), c_void_p(g_mc), arr(g_peers), c_void_p(go_local),
                c_void_p(w_local), c_void_p(w_mc), arr(w_peers), c_void_p(b["vel"]),
                c_void_p(s_local), c_void_p(s_mc), arr(s_peers), c_void_p(f_local), arr(f_peers),
                c_void_p(self.epochs.data_ptr()), c_void_p(self.cta_counter.data_ptr()),
                c_void_p(self.status.data_ptr()), c_float(self.grad_scale), c_int(b["n_ctas"]),
                c_int((1 if self.use_multimem else 0) | (2 if self.shard_master else 0) |
                      (4 if b.get("peer_shadow") else 0) | (8 if (self.split_handshake and not push) else 0) |
                      (16 if push else 0)),
                c_double(self.timeout_s), c_void_p(st_local), arr(st_peers),
                c_void_p(self.comm_stream.cuda_stream))
            _lib.check(rc, "vb_fused_allreduce_gd")
            kernels.LAUNCHES[0] += 2 if self.split_handshake else 1
            done = torch.cuda.Event(enable_timing=prof)
            done.record(self.comm_stream)
            if prof:
                self._prof_events.append((b["channel"], b["bytes"], None, t0, done))
        self._pending_events.append(done)

    def bucket_times(self):
        """[(bucket, parameter bytes, kernel ms)] of the profiled exchange launches since the last call (eager runs
        with ``profile = True``; includes the time the kernel waited for its peers)."""
        import torch
        torch.cuda.synchronize()
        out = [(ch, nbytes, t0.elapsed_time(t1)) for ch, nbytes, _, t0, t1 in self._prof_events]
        self._prof_events = []
        return out

    def allreduce_host(self, grad):
        """Host mode: sum a numpy gradient over the ranks (scaled like the device kernel scales it)."""
        import torch
        if grad is None:
            return None
        t = torch.from_numpy(numpy.ascontiguousarray(grad, dtype=numpy.float32).copy())
        self._dist.all_reduce(t)
        out = t.numpy()
        if self.grad_scale != 1.0:
            out = out * numpy.float32(self.grad_scale)
        return out.reshape(grad.shape)

    def finish_step(self):
        """Join the comm stream into the compute stream and flip the gradient parity."""
        if self.mode == "host":
            return
        import torch
        if self.mode == "fused":
            for b in self.buckets:        # a bucket whose closing unit did not run this step (gated off)
                if b["pending"]:
                    self._launch_bucket(b)
        main = torch.cuda.current_stream()
        for ev in self._pending_events:
            main.wait_event(ev)
        self._pending_events = []
        if self.mode == "fused":
            if self.exchange == "pull":
                self.parity ^= 1      # (push: nobody reads another rank's gradients, one buffer does)
            if self._reserved:
                from ..ops import _lib
                _lib.lib().vb_set_sm_reserve(c_int(0))
                self._reserved = False

    def slice_bounds(self, numel, rank):
        """[lo, hi) of the flat segment that ``rank`` reduces and updates (same arithmetic as the kernel)."""
        chunk = (numel + self.world_size - 1) // self.world_size
        chunk = (chunk + 3) & ~3
        lo = min(numel, chunk * rank)
        return lo, min(numel, lo + chunk)

    def sync_master(self):
        """Collective: reassemble the full fp32 master weights and momentum on every rank from the owners' slices.
        Device-side only: a flag barrier over peer memory, one gather kernel per slab that reads the peers' slices,
        a closing barrier - no NCCL collective and no host synchronisation (the caller's next host read waits on
        the stream as usual).  No-op unless ``shard_master`` left them stale."""
        import torch
        if self.mode != "fused" or not self.master_stale:
            return
        from ..ops import _lib
        L = _lib.lib()
        self.finish_step_events()
        stream = c_void_p(torch.cuda.current_stream().cuda_stream)
        self.device_allreduce(None)            # every rank's last update has completed
        launched = False
        by_slab = {}
        for gd in self.units:
            info = self.unit_info[id(gd)]
            # fp32 master weights, and the momentum each owner alone keeps (a snapshot written by rank 0 must carry
            # all of it: a resumed or re-sharded run continues with the right velocity)
            kinds = (["weights"] if self.shard_master else []) + (["velocity"] if info.get("vel_symmetric") else [])
            for kind in kinds:
                segs = by_slab.setdefault((kind, info["slab"]), [])
                for seg in info["segs"]:
                    if kind == "weights" and (int(seg[5]) & SEG_BCAST_MASTER):
                        continue               # already current on every rank
                    segs.append((int(seg[0]), int(seg[1])))
        for (kind, slab_idx), segs in by_slab.items():
            local, _, peers = self._slab_ptrs(kind, slab_idx)
            peer_arr = (c_void_p * MAX_WORLD)(*[c_void_p(p) for p in peers])
            for k in range(0, len(segs), MAX_SEGS):
                part = segs[k:k + MAX_SEGS]
                seg_arr = (c_longlong * (2 * len(part)))(*[v for sg in part for v in sg])
                rc = L.vb_gather_slices(c_int(self.world_size), c_int(self.rank), c_int(len(part)), seg_arr,
                                        c_void_p(local), peer_arr, c_int(32), stream)
                _lib.check(rc, "vb_gather_slices")
                launched = True
        if launched:
            self.device_allreduce(None)        # nobody updates a slice a peer may still be reading
        self.master_stale = False

    def device_allreduce(self, values, op="sum"):
        """All-reduce a few doubles over the group with ``vb_counter_allreduce`` (flags and slots in symmetric memory);
        ``values`` None = barrier only.  Returns a numpy array (one D2H read: the caller is host logic)."""
        import torch
        from ..ops import _lib
        n = 0 if values is None else len(values)
        if n > COUNTER_SLOT:
            raise ValueError("device_allreduce: %d values exceed the %d-element slot" % (n, COUNTER_SLOT))
        if n:
            self.counter_host[:n] = torch.as_tensor(numpy.asarray(values, dtype=numpy.float64))
            self.counter_in[:n].copy_(self.counter_host[:n], non_blocking=True)
        peer_arr = (c_void_p * MAX_WORLD)(*[c_void_p(p) for p in self.counter_ptrs[1]])
        rc = _lib.lib().vb_counter_allreduce(
            c_int(self.world_size), c_int(self.rank), c_int(n), c_int(1 if op == "max" else 0),
            c_longlong(COUNTER_SLOT), c_void_p(self.counter_in.data_ptr()), c_void_p(self.counter_out.data_ptr()),
            c_void_p(self.counter_ptrs[0]), peer_arr, c_void_p(self.counter_epoch.data_ptr()),
            c_void_p(self.status.data_ptr()), c_double(self.timeout_s),
            c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, "vb_counter_allreduce")
        if not n:
            return None
        out = self.counter_out[:n].cpu().numpy()
        self.check()
        return out

    def finish_step_events(self):
        import torch
        main = torch.cuda.current_stream()
        for ev in self._pending_events:
            main.wait_event(ev)
        self._pending_events = []

    def check(self):
        """Raise if any fused kernel hit its synchronisation timeout (reads one int from the device)."""
        if self.mode == "fused":
            st = int(self.status.item())
            if st > 1000:
                raise RuntimeError("counter exchange timed out waiting for peers (a rank is missing or hung)")
            if st:
                raise RuntimeError("fused all-reduce+GD kernel timed out waiting for peers (channel %d)" % (st - 1))

    # -- epoch-boundary reductions (Decision) -----------------------------------------------------------------------------
    def sum_counters(self, vals, loss):
        import torch
        self.sync_master()   # class boundary: the host (Decision, Snapshotter, plotters) may read the weights now
        if self.mode == "fused":
            out = self.device_allreduce(list(map(float, vals)) + [float(loss)])
        else:
            t = torch.tensor(list(map(float, vals)) + [float(loss)], dtype=torch.float64, device=self._reduce_device())
            self._dist.all_reduce(t)
            out = t.cpu().numpy()
        return numpy.round(out[:-1]).astype(numpy.int64), float(out[-1])

    def max_min(self, mx, mn):
        import torch
        if self.mode == "fused":
            out = self.device_allreduce([mx, -mn], op="max")
        else:
            t = torch.tensor([mx, -mn], dtype=torch.float64, device=self._reduce_device())
            self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX)
            out = t.cpu().numpy()
        return float(out[0]), float(-out[1])

    def sum_array(self, arr, op="sum"):
        """Sum (or max) a small host array over the ranks (confusion matrices, max_err_y_sum); arrays larger than the
        device slot go through the process group."""
        import torch
        a = numpy.asarray(arr)
        if self.mode == "fused" and a.size <= COUNTER_SLOT:
            out = self.device_allreduce(a.reshape(-1).astype(numpy.float64), op=op)
        else:
            t = torch.tensor(a.reshape(-1).astype(numpy.float64), device=self._reduce_device())
            self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX if op == "max" else self._dist.ReduceOp.SUM)
            out = t.cpu().numpy()
        out = out.reshape(a.shape)
        return numpy.round(out).astype(a.dtype) if numpy.issubdtype(a.dtype, numpy.integer) else out.astype(a.dtype)

    def _reduce_device(self):
        return "cpu" if self.mode == "host" else "cuda"

    def barrier(self):
        self._dist.barrier()
