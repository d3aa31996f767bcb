"""Graph-captured training step for the reference's ConvNet (launch_dist.py:9-41, loop :80-92).

One CUDA graph holds the whole iteration: forward kernels (conv+bias+ReLU, pool with argmax),
loss, backward kernels (pool-backward with the ReLU mask folded in, wgrad written straight into
the gradient bucket, dgrad), and the peer-to-peer allreduce fused with 1/W averaging and SGD.
A step is then: two async H2D copies (if the batch is in host memory) + one graph launch.

Buffers are static per batch size; the gradient bucket is DDP's symmetric arena when the model
is wrapped (so the in-graph allreduce kernel reads the peers' buckets over NVLink), or a private
flat buffer for a single un-wrapped GPU.
"""
from __future__ import annotations

import ctypes
import os

import torch

from .. import _lib
from . import functional as NF

_p = NF._p


class _SoloArena:
    """World-size-1 stand-in for DDP's bucket arena (no process group needed)."""

    def __init__(self, params, opt_group, optimizer, device):
        offs, off = [], 0
        for p in params:
            offs.append(off)
            off = (off + p.numel() + 3) // 4 * 4
        self.total, self.offsets = off, offs
        self.g_in = torch.zeros(off, dtype=torch.float32, device=device)
        self.g_avg = torch.zeros(off, dtype=torch.float32, device=device)
        self.p_flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.m_flat = torch.zeros(off, dtype=torch.float32, device=device) if opt_group["momentum"] != 0 else None
        self.states = [torch.zeros(_lib.cuda().pdl_comm_state_bytes(), dtype=torch.uint8, device=device) for _ in range(5)]
        self.state = self.states[0]
        # one device-resident SgdHyper per kernel that applies an update (each clears its own first_step flag):
        # [whole arena, fc1 part, conv3 part, conv1 tail, conv2 part]
        self.hypers = [torch.zeros(8, dtype=torch.float32, device=device) for _ in range(5)]
        self.hyper = self.hypers[0]
        self.hyper_vals = None
        self.first = 1
        with torch.no_grad():
            for p, o in zip(params, offs):
                v = self.p_flat[o:o + p.numel()].view(p.shape)
                v.copy_(p.data)
                p.data = v
                if self.m_flat is not None:
                    mv = self.m_flat[o:o + p.numel()].view(p.shape)
                    old = optimizer.state[p].get("momentum_buffer") if p in optimizer.state else None
                    if old is not None:
                        mv.copy_(old)
                        self.first = 0
                    optimizer.bind_momentum_view(p, mv)
        optimizer.add_reload_listener(self._on_optimizer_reload)

    def _on_optimizer_reload(self, first):
        self.hyper_vals, self.first = None, first

    def refresh_hyper(self, group):
        vals = (float(group["lr"]), float(group["momentum"]), float(group["dampening"]), float(group["weight_decay"]), int(bool(group["nesterov"])))
        if vals != self.hyper_vals:
            h = torch.zeros(8, dtype=torch.float32)
            h[0], h[1], h[2], h[3] = vals[:4]
            hi = h.view(torch.int32)
            hi[4], hi[5] = vals[4], self.first
            for dev_h in self.hypers:
                dev_h.copy_(h)
            self.hyper_vals, self.first = vals, 0

    def launch(self, stream_handle, offset=0, numel=None, which=0):
        """Fused scale+SGD over ``numel`` floats of the arena starting at ``offset`` (default: all of it) -- world size 1, no
        peers.  Parts launched from different streams use different communicator states (retire counters)."""
        L = _lib.cuda()
        n4 = ((self.total if numel is None else numel) + 3) // 4
        ins = (ctypes.c_void_p * 1)(self.g_in.data_ptr() + 4 * offset)
        outs = (ctypes.c_void_p * 1)(self.g_avg.data_ptr() + 4 * offset)
        if which >= len(self.states):
            raise IndexError(which)
        mom = ctypes.c_void_p(self.m_flat.data_ptr() + 4 * offset) if self.m_flat is not None else None
        NF._ok(L.pdl_allreduce_launch(0, 0, 1, 0, n4, ins, outs, None, None, None, None, ctypes.c_void_p(self.p_flat.data_ptr() + 4 * offset), mom,
                                      _p(self.hypers[which]), _p(self.states[which]), L.pdl_allreduce_blocks(n4, 1, 0), 1.0, 30.0,
                                      ctypes.c_void_p(stream_handle)), "fused SGD kernel")


class ConvNetEngine:
    NAMES = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "conv3.weight", "conv3.bias", "fc1.weight", "fc1.bias"]

    @classmethod
    def try_create(cls, model, criterion, optimizer):
        from ..models.convnet import ConvNet
        from ..optim.sgd import SGD
        from ..parallel.ddp import DistributedDataParallel

        ddp = model if isinstance(model, DistributedDataParallel) else None
        module = model.module if ddp is not None else model
        if not isinstance(module, ConvNet) or not isinstance(optimizer, SGD):
            return None
        if not isinstance(criterion, torch.nn.CrossEntropyLoss) or criterion.reduction != "mean" or criterion.weight is not None \
                or criterion.label_smoothing != 0.0:
            return None
        params = dict(module.named_parameters())
        if list(params.keys()) != cls.NAMES or any(p.dtype != torch.float32 or not p.is_cuda or not p.requires_grad for p in params.values()):
            return None
        if len(optimizer.param_groups) != 1 or optimizer.param_groups[0].get("maximize", False):
            return None
        owned = {id(p) for p in optimizer.param_groups[0]["params"]}
        if any(id(p) not in owned for p in params.values()):
            return None
        if ddp is not None and (len(ddp._buckets) != 1 or not getattr(ddp._backend, "supports_fused_sgd", False) or not ddp.fused_optimizer):
            return None
        if ddp is None:
            return cls(module, ddp, optimizer)
        # one engine per (wrapped model, optimizer): an explicit TrainStep and the transparent six-call path share graphs,
        # buffers and the reserved signal-pad channels instead of claiming them twice
        cache = ddp.__dict__.setdefault("_convnet_engines", {})
        engine = cache.get(id(optimizer))
        if engine is None or engine.optimizer is not optimizer:
            engine = cache[id(optimizer)] = cls(module, ddp, optimizer)
        return engine

    def __init__(self, module, ddp, optimizer):
        self.module, self.ddp, self.optimizer = module, ddp, optimizer
        self.params = [dict(module.named_parameters())[n] for n in self.NAMES]
        self.device = self.params[0].device
        self.L = _lib.cuda()
        if ddp is not None:
            if not ddp._fused_checked:
                ddp._resolve_fused_optimizers()
            bk = ddp._buckets[0]
            if bk.fused_group is None:
                raise RuntimeError("TrainStep: the optimizer could not be fused into the bucket allreduce")
            ddp._rebuilt = True  # the captured graph bakes in arena offsets: freeze the bucket layout
            self.backend, self.bucket = ddp._backend, bk
            index_of = {id(p): i for i, p in enumerate(ddp._params)}
            self.g_views = [ddp._in_views[index_of[id(p)]] for p in self.params]
            self.avg_views = [ddp._avg_views[index_of[id(p)]] for p in self.params]
            self.world = ddp._world
        else:
            self.arena = _SoloArena(self.params, self.group, optimizer, self.device)
            a = self.arena
            self.g_views = [a.g_in[o:o + p.numel()].view(p.shape) for p, o in zip(self.params, a.offsets)]
            self.avg_views = [a.g_avg[o:o + p.numel()].view(p.shape) for p, o in zip(self.params, a.offsets)]
            self.world = 1
        for p, v in zip(self.params, self.avg_views):
            p.grad = v
        self._setup_split()
        if ddp is not None:
            # resolve variant / CTA count / push inbox of the whole-bucket kernel NOW (strict-fp32 schedule): the inbox is a
            # collective allocation and must not happen while a graph is being captured
            self.backend.bucket_launch_args(self.bucket, self.world, ddp.algorithm)
        optimizer.add_reload_listener(self._on_optimizer_reload)
        self.graphs = {}
        self.kernel_count = 0
        bprio = -1 if os.environ.get("PDL_CONVNET_PRIORITY", "0") == "2" else 0
        self._branch = torch.cuda.Stream(device=self.device, priority=bprio)   # off-critical-path kernels (see _record_tc): pack, fc1 wgrad
        self._branch2 = torch.cuda.Stream(device=self.device, priority=bprio)  # conv3 / conv2 weight gradients
        self._capture_stream = torch.cuda.Stream(device=self.device, priority=-1)  # the main chain: high priority (see _build)
        self._branch3 = torch.cuda.Stream(device=self.device, priority=-1)  # conv2's wgrad chain: high priority (see _record_tc_impl)
        self._branch4 = torch.cuda.Stream(device=self.device)               # conv3's gradient exchange
        self._ring = self._d2h = None  # pinned loss ring + its copy stream, created by the first step(pending=True)

    def _on_optimizer_reload(self, first):
        self._first_cleared = False  # the restored state decides the first-step flags again (see _clear_first_flags)

    @property
    def group(self):
        return self.optimizer.param_groups[0]  # looked up every time: load_state_dict() replaces the dict

    # ------------------------------------------------------------------ two-part gradient exchange
    def _setup_split(self):
        """The arena is laid out in gradient-arrival order -- fc1 | conv3 | conv2 | conv1 -- so each layer is a contiguous range.
        Instead of ONE exchange after the last gradient (fully exposed: nothing is left to overlap it with), the step
        exchanges four parts, each the moment it is final, each fused with its SGD update:
          * fc1    an allreduce+SGD kernel on branch stream A, right behind fc1's weight-gradient kernel;
          * conv3  an allreduce+SGD kernel on a stream of its own behind conv3's weight gradient (hidden: half the backward
                   pass is still to come);
          * conv2  INSIDE the second pass of its split-K weight gradient (csrc/cuda/tc_wgrad.cu: every CTA reduces its partial
                   sums, pushes the result to the peers, one flag round, reduce, SGD) -- no separate collective kernel on the
                   branch that ends the step;
          * conv1  (832 floats, the LAST gradient) inside the conv1-wgrad kernel (csrc/cuda/convnet_bwd.cu).
        Kernels on different streams can be in flight together, so every part owns its flags: a reserved signal-pad channel
        of the arena's segment (fc1, conv3) or a small segment of its own with inbox + pad + state (conv2, conv1)."""
        from ..parallel.ddp import _Bucket

        self.split = None
        if os.environ.get("PDL_CONVNET_SPLIT", "1") == "0":
            return
        flat_in = self.backend.g_in if self.ddp is not None else self.arena.g_in
        esz = flat_in.element_size()
        off = lambda i: (self.g_views[i].data_ptr() - flat_in.data_ptr()) // esz  # noqa: E731
        base = self.bucket.offset if self.ddp is not None else 0
        total = (self.bucket.offset + self.bucket.numel) if self.ddp is not None else self.arena.total
        off_fcb, off_c3b, off_c2b, off_c2w, off_c1b, off_c1w = off(7), off(5), off(3), off(2), off(1), off(0)
        ok = off_fcb == base and base < off_c3b < off_c2b and off_c2b + 64 == off_c2w and off_c2w + 18432 == off_c1b \
            and off_c1b + 32 == off_c1w and off_c1w + 800 == total and off_c3b % 4 == 0 and off_c2b % 4 == 0 and off_c1b % 4 == 0 \
            and off(6) < off_c3b and off_c3b <= off(4) < off_c2b
        if not ok:
            return  # unexpected layout (e.g. a user-forced bucket order): keep the single whole-bucket kernel
        # conv2's exchange inside its second pass is OPT-IN: measured on B200 at batch 100 it is 3 us/step slower on one GPU
        # (104.4 vs 101.4 us: one output per thread against the float4 allreduce kernel) and within noise on two (116.2 vs 117.0)
        fuse_c2 = os.environ.get("PDL_CONVNET_FUSE_C2", "0") == "1"
        # conv3 and conv2 are exchanged separately (PDL_CONVNET_JOIN_C32=1 restores one kernel over both): conv3's part leaves
        # from a stream of its own as soon as its weight gradient is final, so what is left for the end of the step is conv2's
        # 74 KB instead of 369 KB
        join_c32 = os.environ.get("PDL_CONVNET_JOIN_C32", "0") == "1" and not fuse_c2
        split = {"fc": (base, off_c3b - base), "c3": (off_c3b, (off_c1b if join_c32 else off_c2b) - off_c3b), "c2": (off_c2b, off_c1b - off_c2b),
                 "join_c32": join_c32,
                 "fuse_c2": fuse_c2, "tail": {"inbox": None, "pads": None, "state": None}, "c2x": {"inbox": None, "pads": None, "state": None}}
        L = self.L
        if self.ddp is not None:
            parts = {}
            for name, index, (o, n) in (("fc", 1000, split["fc"]), ("c3", 1002, split["c3"]), ("tail", 1001, (off_c1b, 832)), ("c2", 1003, split["c2"])):
                bk = _Bucket(index, [], o, n)
                bk.fused_group, bk.first_flag = self.bucket.fused_group, self.bucket.first_flag
                parts[name] = bk
            parts["fc"].channel = self.backend.reserve_channel()  # launched from branch streams: never share flags with the
            parts["c3"].channel = self.backend.reserve_channel()  # communication stream's buckets or with each other
            parts["c2"].channel = self.backend.reserve_channel()
            split["parts"] = parts
            pg = self.backend.pg
            split["c2x"]["state"] = pg.new_comm_state()  # also at world size 1: the fused second pass counts its CTAs in it
            if self.world > 1:
                pad = (L.pdl_signal_pad_bytes() + 4095) // 4096 * 4096
                for key, nbytes in (("tail", int(L.pdl_conv1_tail_inbox_bytes(self.world))),
                                    ("c2x", 4 * int(L.pdl_wgrad_finish_comm_inbox_floats(self.world, 32, 64, 3, 3)))):
                    seg = pg.new_segment(pad + nbytes, want_multicast=False)
                    d = split[key]
                    d["seg"] = seg
                    if d["state"] is None:
                        d["state"] = pg.new_comm_state()
                    d["pads"] = (ctypes.c_void_p * self.world)(*[seg.ptr(r) for r in range(self.world)])
                    d["inbox"] = (ctypes.c_void_p * self.world)(*[seg.ptr(r, pad) for r in range(self.world)])
                for name in ("fc", "c3", "c2"):  # resolve variant / CTAs / inbox now: collective allocations stay out of graph capture
                    self.backend.bucket_launch_args(parts[name], self.world, self.ddp.algorithm)
        else:
            split["c2x"]["state"] = self.arena.states[4]
        self.split = split

    def _part_hyper(self, name):
        if self.ddp is not None:
            return self.backend._hyper_for(self.split["parts"][name])
        return self.arena.hypers[{"fc": 1, "c3": 2, "tail": 3, "c2": 4}[name]]

    def _tail_hyper(self):
        return self._part_hyper("tail")

    def _launch_part(self, name, stream_handle):
        """Exchange + SGD of one contiguous part of the arena ("fc", "c3" or "c2") on the given stream."""
        if self.ddp is not None:
            self.backend.launch_bucket_kernel(self.split["parts"][name], self.world, self.ddp.algorithm, stream_handle, fused=True)
        else:
            o, n = self.split[name]
            self.arena.launch(stream_handle, offset=o, numel=n, which={"fc": 1, "c3": 2, "c2": 4}[name])

    def _launch_conv2_finish_comm(self, b, N, st):
        """Second pass of conv2's weight gradient + cross-GPU average + SGD in one kernel (the conv2 slice of the arena)."""
        sp, x = self.split, self.split["c2x"]
        o, n = sp["c2"]
        flat = (lambda t: ctypes.c_void_p(t.data_ptr() + 4 * o))
        if self.ddp is not None:
            bk = self.backend
            g_in, g_avg, p_flat, m_flat = bk.g_in, bk.g_avg, bk.p_flat, bk.m_flat
            rank, timeout = bk.pg.rank(), bk.pg.timeout_s
        else:
            a = self.arena
            g_in, g_avg, p_flat, m_flat = a.g_in, a.g_avg, a.p_flat, a.m_flat
            rank, timeout = 0, 30.0
        mom = flat(m_flat) if (self.group["momentum"] != 0 and m_flat is not None) else None
        NF._ok(self.L.pdl_tc_conv2d_wgrad_finish_comm(_p(b["wg_scratch"]), _p(b["db2part"]), self.db2_blocks, N, 13, 13, 32, 64, 3, 3, 0,
                                                      flat(g_in), flat(g_avg), flat(p_flat), mom, _p(self._part_hyper("c2")), rank, self.world, 0,
                                                      x["inbox"], x["pads"], _p(x["state"]), 1.0 / self.world, timeout, st),
               "conv2 wgrad second pass + gradient exchange + SGD")

    def _launch_conv1_wgrad_comm(self, b, N, st):
        """conv1 weight gradient + cross-GPU average + SGD in one kernel (the tail of the arena)."""
        sp = self.split["tail"]
        w1, b1 = self.params[0], self.params[1]
        gw1, gb1 = self.g_views[0], self.g_views[1]
        aw1, ab1 = self.avg_views[0], self.avg_views[1]
        mom = self.group["momentum"] != 0
        mw = _p(self.optimizer.state[w1]["momentum_buffer"]) if mom else None
        mb = _p(self.optimizer.state[b1]["momentum_buffer"]) if mom else None
        rank = self.backend.pg.rank() if self.ddp is not None else 0
        timeout = self.backend.pg.timeout_s if self.ddp is not None else 30.0
        NF._ok(self.L.pdl_conv1_wgrad_fused_comm(_p(b["x"]), _p(b["da1"]), _p(b["i1"]), _p(b["a1"]), _p(gw1), _p(gb1), _p(b["c1part"]), N,
                                                 _p(aw1), _p(ab1), _p(w1), _p(b1), mw, mb, _p(self._tail_hyper()), rank, self.world, sp["inbox"], sp["pads"],
                                                 _p(sp["state"]) if sp["state"] is not None else None, 1.0 / self.world, timeout, st),
               "conv1 wgrad + gradient exchange + SGD")

    # ------------------------------------------------------------------ buffers + capture
    def _build(self, N):
        dev, f32 = self.device, torch.float32
        e = lambda *s, dt=f32: torch.empty(s, dtype=dt, device=dev)  # noqa: E731
        b = {
            "x": e(N, 28, 28, 1), "y": e(N, dt=torch.int64),
            "z1": e(N, 26, 26, 32), "a1": e(N, 13, 13, 32), "i1": e(N, 13, 13, 32, dt=torch.uint8),
            "z2": e(N, 11, 11, 64), "a2": e(N, 10, 10, 64), "i2": e(N, 10, 10, 64, dt=torch.uint8),
            "z3": e(N, 8, 8, 128), "a3": e(N, 4, 4, 128), "i3": e(N, 4, 4, 128, dt=torch.uint8),
            "flat": e(N, 2048), "logits": e(N, 10), "row_loss": e(N), "loss": torch.zeros((), dtype=f32, device=dev), "dlogits": e(N, 10),
            "dflat": e(N, 2048), "da3": e(N, 4, 4, 128), "dz3": e(N, 8, 8, 128), "da2": e(N, 10, 10, 64), "dz2": e(N, 11, 11, 64),
            "da1": e(N, 13, 13, 32), "dz1": e(N, 26, 26, 32),
        }
        shapes = [(N, 28, 28, 1, 32, 5, 5, 1, 1), (N, 13, 13, 32, 64, 3, 3, 1, 0), (N, 10, 10, 64, 128, 3, 3, 1, 0)]
        nbytes = max(self.L.pdl_conv2d_wgrad_scratch_bytes(*s) for s in shapes)
        b["scratch"] = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
        from . import tc_conv

        self.use_tc = tc_conv.enabled()
        if self.use_tc:  # tensor-core path: packed weights, bias partials, split-K scratch
            L = self.L
            blocks = L.pdl_bwd_blocks(N)
            # measured on B200 at batch 100: pool2-backward in conv3's dgrad epilogue is neutral (132.7 vs 132.6 us/step) and costs
            # 5 % at batch 4096, so it is opt-in
            self.fuse_pool2 = os.environ.get("PDL_CONVNET_FUSE_POOL2", "0") == "1" and L.pdl_tc_dgrad_poolbwd_tiles(N, 10, 10) > 0
            pool2_blocks = L.pdl_tc_dgrad_poolbwd_tiles(N, 10, 10) if self.fuse_pool2 else L.pdl_pool2_bwd_blocks(N, 11, 11, 64)
            wg = max(L.pdl_tc_wgrad_scratch_bytes(N, 13, 13, 32, 64, 3, 3, 0), L.pdl_tc_wgrad_scratch_bytes(N, 10, 10, 64, 128, 3, 3, 0))
            b.update({
                "p2f": e(9, 64, 32), "p2b": e(9, 32, 64), "p3f": e(9, 128, 64), "p3b": e(9, 64, 128), "wfp": e(10, 2048),
                "fc_scratch": e(int(L.pdl_fc_bwd_scratch_floats(N))), "db3part": e(N, 128), "db2part": e(pool2_blocks, 64),
                "c1part": e(blocks, 832), "wg_scratch": torch.empty(int(wg), dtype=torch.uint8, device=dev),
                "err": tc_conv.error_flag(dev),
            })
            self.db2_blocks = pool2_blocks
        b["x"].zero_()
        b["y"].zero_()
        entry = {"buf": b, "graph": None, "kernels": 0, "tc": self.use_tc}
        # warm-up run outside capture (also validates every launch), then capture
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            self._record(b, N, side.cuda_stream, reduce=False)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        # Kernel nodes inherit the priority of the stream they were captured on.  The main chain is the step's critical path
        # and the branch streams only carry work that has slack (weight gradients, their exchange), so the chain is captured
        # on a HIGH-priority stream: whenever CTAs of both are waiting for SMs -- the tcgen05 kernels take a whole SM's shared
        # memory each, so concurrent kernels queue for SMs rather than share them -- the chain's go first.
        prio = os.environ.get("PDL_CONVNET_PRIORITY", "0") == "1"
        cap = dict(stream=self._capture_stream) if prio else {}
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, **cap):
            entry["kernels"] = self._record(b, N, torch.cuda.current_stream(dev).cuda_stream, reduce=True)
        entry["graph"] = g
        # Slot 1: the same iteration captured over a second set of input/loss buffers (everything else is shared --
        # replays are ordered on one stream).  Host batches alternate between the two, so the H2D copy of batch i+1
        # lands while step i still reads batch i, and the loss of step i can leave for the host while step i+1 runs.
        b1 = dict(b)
        b1["x"], b1["y"], b1["loss"] = torch.zeros_like(b["x"]), torch.zeros_like(b["y"]), torch.zeros_like(b["loss"])
        torch.cuda.synchronize(dev)
        g1 = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g1, **cap):
            self._record(b1, N, torch.cuda.current_stream(dev).cuda_stream, reduce=True)
        from ..parallel.pipeline import InputStager

        entry["slots"] = [{"buf": b, "graph": g, "read": None}, {"buf": b1, "graph": g1, "read": None}]
        entry["stager"] = InputStager(dev, zones=[(b["x"], b["y"]), (b1["x"], b1["y"])])
        return entry

    def _record(self, b, N, stream_handle, reduce):
        if self.use_tc and self.split is not None:
            return self._record_tc(b, N, stream_handle, reduce)  # the gradient exchange is part of the schedule
        k = self._record_tc(b, N, stream_handle) if self.use_tc else self._record_simt(b, N, stream_handle)
        # gradient averaging + SGD (one kernel; over NVLink when world > 1)
        if reduce:
            if self.ddp is not None:
                self.backend.launch_bucket_kernel(self.bucket, self.world, self.ddp.algorithm, stream_handle, fused=True)
            else:
                self.arena.launch(stream_handle)
            k += 1
        return k

    def _record_tc(self, b, N, stream_handle, reduce=False):
        # the two weight gradients run beside the main chain: a third of a wave of CTAs each (see pdl_tc_wgrad_max_ctas)
        # -- only while the step is latency-bound; at large batches the weight gradients need the whole machine
        cap = int(os.environ.get("PDL_CONVNET_WG_CTAS", "48")) if N <= 256 else 0
        old_cap = self.L.pdl_tc_wgrad_max_ctas(cap)
        try:
            return self._record_tc_impl(b, N, stream_handle, reduce)
        finally:
            self.L.pdl_tc_wgrad_max_ctas(old_cap)

    def _record_tc_impl(self, b, N, stream_handle, reduce=False):
        """Tensor-core path: 15 launches (tcgen05 conv2/conv3 fwd+dgrad+wgrad, fused CUDA-core kernels for the rest).

        At batch 100 every kernel is a fraction of one wave and latency-bound, so the step time is the length of the
        dependency chain.  The weight gradients (+ their split reductions) and the weight packing are not on that
        chain: they are recorded on a branch stream (fork/join through events, which capture turns into graph edges)
        and overlap the dgrad / pool-backward kernels of the main chain."""
        L, ok, st = self.L, NF._ok, ctypes.c_void_p(stream_handle)
        w1, b1, w2, b2, w3, b3, wf, bf = self.params
        gw1, gb1, gw2, gb2, gw3, gb3, gwf, gbf = self.g_views
        err = _p(b["err"])
        k = 0
        main = torch.cuda.ExternalStream(stream_handle, device=self.device)
        fork = os.environ.get("PDL_CONVNET_FORK", "1") != "0"
        branch = self._branch if fork else main
        # the conv weight gradients do not depend on fc1's: a second branch lets conv3's wgrad start as soon as dz3 exists
        # instead of queueing behind fc1's 20 us weight-gradient kernel (timeline: profiles/convnet_timeline_w2.txt)
        branch2 = self._branch2 if fork and os.environ.get("PDL_CONVNET_BRANCHES", "2") != "1" else branch
        sb = ctypes.c_void_p(branch.cuda_stream)
        sb2 = ctypes.c_void_p(branch2.cuda_stream)
        exchange = reduce and self.split is not None
        joined = [branch2]
        branch.wait_stream(main)
        ok(L.pdl_convnet_pack(_p(w2), _p(w3), _p(wf), _p(b["p2f"]), _p(b["p2b"]), _p(b["p3f"]), _p(b["p3b"]), _p(b["wfp"]), sb), "weight pack"); k += 1
        # forward: conv+bias+ReLU+pool fused per layer
        ok(L.pdl_conv1_fused_fwd(_p(b["x"]), _p(w1), _p(b1), _p(b["a1"]), _p(b["i1"]), N, st), "conv1"); k += 1
        main.wait_stream(branch)
        ok(L.pdl_tc_conv2d_fwd(_p(b["a1"]), _p(b["p2f"]), _p(b2), _p(b["a2"]), _p(b["i2"]), N, 13, 13, 32, 64, 3, 3, 0, 1, 1, err, st), "conv2"); k += 1
        ok(L.pdl_tc_conv2d_fwd(_p(b["a2"]), _p(b["p3f"]), _p(b3), _p(b["a3"]), _p(b["i3"]), N, 10, 10, 64, 128, 3, 3, 0, 1, 2, err, st), "conv3"); k += 1
        if fork:
            # fc1's weight gradient needs every row's dlogits and is only consumed by the final allreduce: branch.
            # dz3 feeds the dgrad chain on main.  A row's dz3 only needs that row's dlogits, so the fc1+loss kernel can compute
            # it too (PDL_CONVNET_FUSE_FC=1, one CTA per sample) -- measured SLOWER on B200 (132.7 vs 123.9 us/step at batch
            # 100: the two latency chains serialise inside each CTA), hence off by default.
            if os.environ.get("PDL_CONVNET_FUSE_FC", "0") == "1":
                ok(L.pdl_fc_ce_bwd_fused(_p(b["a3"]), _p(b["wfp"]), _p(bf), _p(b["y"]), _p(b["logits"]), _p(b["row_loss"]), _p(b["loss"]),
                                         _p(b["dlogits"]), _p(b["i3"]), _p(b["dz3"]), _p(b["db3part"]), N, st), "fc1 + loss + fc1 dgrad + pool3 backward"); k += 1
                part2 = False
            else:
                ok(L.pdl_fc_ce_fused(_p(b["a3"]), _p(b["wfp"]), _p(bf), _p(b["y"]), _p(b["logits"]), _p(b["row_loss"]), _p(b["loss"]), _p(b["dlogits"]),
                                     N, st), "fc1 + loss"); k += 1
                part2 = True
            branch.wait_stream(main)
            ok(L.pdl_fc_bwd_pool3_part(1, _p(b["a3"]), _p(b["i3"]), _p(b["dlogits"]), _p(b["wfp"]), _p(gwf), _p(gbf), _p(b["fc_scratch"]), _p(b["dz3"]),
                                       _p(b["db3part"]), N, sb), "fc1 weight gradient"); k += 1
            if exchange:
                self._launch_part("fc", branch.cuda_stream); k += 1
            if part2:
                ok(L.pdl_fc_bwd_pool3_part(2, _p(b["a3"]), _p(b["i3"]), _p(b["dlogits"]), _p(b["wfp"]), _p(gwf), _p(gbf), _p(b["fc_scratch"]),
                                           _p(b["dz3"]), _p(b["db3part"]), N, st), "fc1 dgrad + pool3 backward"); k += 1
        else:
            ok(L.pdl_fc_ce_fused(_p(b["a3"]), _p(b["wfp"]), _p(bf), _p(b["y"]), _p(b["logits"]), _p(b["row_loss"]), _p(b["loss"]), _p(b["dlogits"]), N, st),
               "fc1 + loss"); k += 1
            ok(L.pdl_fc_bwd_pool3_fused(_p(b["a3"]), _p(b["i3"]), _p(b["dlogits"]), _p(b["wfp"]), _p(gwf), _p(gbf), _p(b["fc_scratch"]), _p(b["dz3"]),
                                        _p(b["db3part"]), N, st), "fc1 backward + pool3 backward"); k += 1
        if exchange and not fork:
            self._launch_part("fc", branch.cuda_stream); k += 1
        branch2.wait_stream(main)  # dz3 / db3part are ready
        ok(L.pdl_tc_conv2d_wgrad(_p(b["a2"]), _p(b["dz3"]), _p(gw3), _p(gb3), _p(b["db3part"]), N, _p(b["wg_scratch"]), N, 10, 10, 64, 128, 3, 3, 0,
                                 err, sb2), "conv3 wgrad"); k += 2
        fuse_c2 = exchange and self.split["fuse_c2"] and fork
        own_c3 = exchange and fork and not self.split["join_c32"]
        if own_c3:
            # conv3's gradient is final: its exchange + SGD goes to a stream of its own (half of the backward pass is still to
            # come, so it is hidden), leaving this branch free for conv2's weight gradient the moment dz2 exists
            self._branch4.wait_stream(branch2)
            joined.append(self._branch4)
            self._launch_part("c3", self._branch4.cuda_stream); k += 1
        if self.fuse_pool2:  # conv3 dgrad with pool2's backward + ReLU mask + bias partials in its epilogue
            ok(L.pdl_tc_conv2d_dgrad_poolbwd(_p(b["dz3"]), _p(b["p3b"]), _p(b["i2"]), _p(b["a2"]), _p(b["dz2"]), _p(b["db2part"]), N, 10, 10, 64, 128,
                                             3, 3, 0, 1, err, st), "conv3 dgrad + pool2 backward"); k += 1
        else:
            ok(L.pdl_tc_conv2d_dgrad(_p(b["dz3"]), _p(b["p3b"]), _p(b["da2"]), N, 10, 10, 64, 128, 3, 3, 0, err, st), "conv3 dgrad"); k += 1
            ok(L.pdl_pool2_bwd_dbias(_p(b["da2"]), _p(b["i2"]), _p(b["a2"]), _p(b["dz2"]), _p(b["db2part"]), N, 11, 11, 64, 1, st), "pool2 backward"); k += 1
        if fork and N <= 256 and os.environ.get("PDL_CONVNET_WG2_PRIO", "1") == "1":
            # conv2's weight gradient and its dgrad become ready together; on a HIGH-priority stream the (48-CTA) weight gradient
            # takes its SMs at once instead of queueing behind the dgrad's 133 CTAs -- it is what the end of the step waits
            # for (measured: 108.8 -> 104.5 us/step, 103.8 with the dgrad bounded to two tiles per CTA)
            self._branch3.wait_stream(branch2)
            branch2 = self._branch3
            joined.append(branch2)
            sb2 = ctypes.c_void_p(branch2.cuda_stream)
        branch2.wait_stream(main)  # dz2 / db2part are ready (the branch is also past conv3's wgrad: wg_scratch is free again)
        if fuse_c2:
            ok(L.pdl_tc_conv2d_wgrad_gemm(_p(b["a1"]), _p(b["dz2"]), _p(b["wg_scratch"]), N, 13, 13, 32, 64, 3, 3, 0, err, sb2), "conv2 wgrad gemm"); k += 1
            self._launch_conv2_finish_comm(b, N, sb2); k += 1  # second pass + gradient exchange + SGD: one kernel
        else:
            ok(L.pdl_tc_conv2d_wgrad(_p(b["a1"]), _p(b["dz2"]), _p(gw2), _p(gb2), _p(b["db2part"]), self.db2_blocks, _p(b["wg_scratch"]), N, 13, 13, 32, 64,
                                     3, 3, 0, err, sb2), "conv2 wgrad"); k += 2
            if exchange:
                # conv2's gradient (and conv3's, when the two are joined or there is no fork) is final on this branch:
                # exchange + SGD here, behind the two kernels the main chain still has to run
                if not own_c3:
                    self._launch_part("c3", branch2.cuda_stream); k += 1
                if not self.split["join_c32"]:
                    self._launch_part("c2", branch2.cuda_stream); k += 1
        d2cap = int(os.environ.get("PDL_CONVNET_DGRAD2_CTAS", "100")) if N <= 256 else 0
        old = L.pdl_tc_conv_max_ctas(d2cap)  # conv2's dgrad and wgrad become ready together: leave the wgrad some SMs
        ok(L.pdl_tc_conv2d_dgrad(_p(b["dz2"]), _p(b["p2b"]), _p(b["da1"]), N, 13, 13, 32, 64, 3, 3, 0, err, st), "conv2 dgrad"); k += 1
        L.pdl_tc_conv_max_ctas(old)
        if reduce and self.split is not None:
            self._launch_conv1_wgrad_comm(b, N, st); k += 1
        else:
            ok(L.pdl_conv1_wgrad_fused(_p(b["x"]), _p(b["da1"]), _p(b["i1"]), _p(b["a1"]), _p(gw1), _p(gb1), _p(b["c1part"]), N, st), "conv1 wgrad"); k += 1
        main.wait_stream(branch)  # join: every gradient is in the bucket before the allreduce / the step ends
        for other in joined:  # every stream that received work in this capture joins the main chain
            if other is not branch and other is not main:
                main.wait_stream(other)
        return k

    def _record_simt(self, b, N, stream_handle):
        """Strict-fp32 CUDA-core path (used when TF32 convolutions are disabled)."""
        L, ok, st = self.L, NF._ok, ctypes.c_void_p(stream_handle)
        w1, b1, w2, b2, w3, b3, wf, bf = self.params
        gw1, gb1, gw2, gb2, gw3, gb3, gwf, gbf = self.g_views
        k = 0
        # forward
        ok(L.pdl_conv2d_fwd_simt(_p(b["x"]), _p(w1), _p(b1), _p(b["z1"]), N, 28, 28, 1, 32, 5, 5, 1, 1, 1, st), "conv1"); k += 1
        ok(L.pdl_maxpool_fwd(_p(b["z1"]), _p(b["a1"]), _p(b["i1"]), N, 26, 26, 32, 2, 2, 0, st), "pool1"); k += 1
        ok(L.pdl_conv2d_fwd_simt(_p(b["a1"]), _p(w2), _p(b2), _p(b["z2"]), N, 13, 13, 32, 64, 3, 3, 1, 0, 1, st), "conv2"); k += 1
        ok(L.pdl_maxpool_fwd(_p(b["z2"]), _p(b["a2"]), _p(b["i2"]), N, 11, 11, 64, 2, 1, 0, st), "pool2"); k += 1
        ok(L.pdl_conv2d_fwd_simt(_p(b["a2"]), _p(w3), _p(b3), _p(b["z3"]), N, 10, 10, 64, 128, 3, 3, 1, 0, 1, st), "conv3"); k += 1
        ok(L.pdl_maxpool_fwd(_p(b["z3"]), _p(b["a3"]), _p(b["i3"]), N, 8, 8, 128, 2, 2, 0, st), "pool3"); k += 1
        ok(L.pdl_permute_nhwc_nchw(_p(b["a3"]), _p(b["flat"]), N, 16, 128, 1, st), "flatten"); k += 1
        ok(L.pdl_linear_fwd(_p(b["flat"]), _p(wf), _p(bf), _p(b["logits"]), N, 10, 2048, st), "fc1"); k += 1
        ok(L.pdl_cross_entropy(_p(b["logits"]), _p(b["y"]), N, 10, _p(b["row_loss"]), _p(b["loss"]), _p(b["dlogits"]), None, st), "loss"); k += 1
        # backward
        ok(L.pdl_linear_wgrad(_p(b["flat"]), _p(b["dlogits"]), _p(gwf), _p(gbf), N, 10, 2048, st), "fc1 wgrad"); k += 1
        ok(L.pdl_linear_dgrad(_p(b["dlogits"]), _p(wf), _p(b["dflat"]), N, 10, 2048, st), "fc1 dgrad"); k += 1
        ok(L.pdl_permute_nhwc_nchw(_p(b["dflat"]), _p(b["da3"]), N, 16, 128, 0, st), "unflatten"); k += 1
        ok(L.pdl_maxpool_bwd(_p(b["da3"]), _p(b["i3"]), _p(b["a3"]), _p(b["dz3"]), N, 8, 8, 128, 2, 2, 0, 1, st), "pool3 bwd"); k += 1
        ok(L.pdl_conv2d_wgrad_simt(_p(b["a2"]), _p(b["dz3"]), _p(gw3), _p(gb3), _p(b["scratch"]), N, 10, 10, 64, 128, 3, 3, 1, 0, st), "conv3 wgrad"); k += 2
        ok(L.pdl_conv2d_dgrad_simt(_p(b["dz3"]), _p(w3), _p(b["da2"]), N, 10, 10, 64, 128, 3, 3, 1, 0, st), "conv3 dgrad"); k += 1
        ok(L.pdl_maxpool_bwd(_p(b["da2"]), _p(b["i2"]), _p(b["a2"]), _p(b["dz2"]), N, 11, 11, 64, 2, 1, 0, 1, st), "pool2 bwd"); k += 1
        ok(L.pdl_conv2d_wgrad_simt(_p(b["a1"]), _p(b["dz2"]), _p(gw2), _p(gb2), _p(b["scratch"]), N, 13, 13, 32, 64, 3, 3, 1, 0, st), "conv2 wgrad"); k += 2
        ok(L.pdl_conv2d_dgrad_simt(_p(b["dz2"]), _p(w2), _p(b["da1"]), N, 13, 13, 32, 64, 3, 3, 1, 0, st), "conv2 dgrad"); k += 1
        ok(L.pdl_maxpool_bwd(_p(b["da1"]), _p(b["i1"]), _p(b["a1"]), _p(b["dz1"]), N, 26, 26, 32, 2, 2, 0, 1, st), "pool1 bwd"); k += 1
        ok(L.pdl_conv2d_wgrad_simt(_p(b["x"]), _p(b["dz1"]), _p(gw1), _p(gb1), _p(b["scratch"]), N, 28, 28, 1, 32, 5, 5, 1, 1, st), "conv1 wgrad"); k += 2
        return k

    # ------------------------------------------------------------------ stepping
    def _refresh_hyper(self):
        if self.ddp is not None:
            self.backend._hyper_for(self.bucket)
            if self.split is not None:
                for bk in self.split["parts"].values():
                    self.backend._hyper_for(bk)
        else:
            self.arena.refresh_hyper(self.group)

    def _clear_first_flags(self):
        """After the first real step every SgdHyper copy -- also the ones whose kernel did not run in this configuration --
        must say "momentum history exists" (stream-ordered behind the step)."""
        if self.ddp is not None:
            tensors = [e["dev"] for e in self.backend._hyper.values()]
        else:
            tensors = self.arena.hypers
        for t in tensors:
            t.view(torch.int32)[5:6].zero_()
        self._first_cleared = True

    def step(self, images, labels, pending=False):
        """One iteration.  Returns the loss as a device scalar, or (``pending=True``) as a ``PendingLoss`` whose value is
        copied to pinned host memory behind the step.  Host batches are staged through two alternating input slots."""
        N = images.shape[0]
        from . import tc_conv

        key = (N, tc_conv.enabled())
        entry = self.graphs.get(key)
        if entry is None:
            if tuple(images.shape[1:]) != (1, 28, 28):
                raise ValueError("ConvNet expects N x 1 x 28 x 28 inputs, got %s" % (tuple(images.shape),))
            self._refresh_hyper()
            entry = self.graphs[key] = self._build(N)
        self._refresh_hyper()
        if self.ddp is not None:
            self.backend.pg.poll_errors()  # a lost peer reported by an earlier replay's allreduce kernel raises here (one host load)
        if entry["tc"]:
            tc_conv.check_errors(self.device)  # ... and so does a tensor-core kernel of an earlier replay that gave up on its pipeline
        main = torch.cuda.current_stream(self.device)
        stager = entry["stager"]
        if not images.is_cuda and not labels.is_cuda:
            _, _, ready, slot = stager.stage(images, labels)  # H2D on the copy stream, overlapping the step in flight
            main.wait_event(ready)
        else:
            slot = 0
            b0 = entry["buf"]
            main.wait_event(stager.ready[0])  # a staged copy into this slot, if any, is ordered before this one
            from ..parallel.pipeline import refresh_inputs

            refresh_inputs(b0["x"], b0["y"], images, labels)
        s = entry["slots"][slot]
        if s["read"] is not None:  # this slot's previous loss must have left for the host before it is overwritten
            main.wait_event(s["read"])
            s["read"] = None
        s["graph"].replay()
        if not getattr(self, "_first_cleared", False):
            self._clear_first_flags()
        stager.consumed(slot, main)
        self.kernel_count = entry["kernels"]
        self.last_key = key
        self.serial = getattr(self, "serial", 0) + 1
        self._last_loss = s["buf"]["loss"]
        if self.ddp is not None:
            self.ddp._iteration += 1
        if pending:
            if self._ring is None:
                from ..parallel.pipeline import LossRing

                self._ring, self._d2h = LossRing(self.device), torch.cuda.Stream(device=self.device)
            handle, s["read"] = self._ring.publish(s["buf"]["loss"], self._d2h, after=stager.free[slot])
            return handle
        return s["buf"]["loss"]

    def last_loss(self):
        """Device scalar holding the most recent step's loss (static buffer: the next step on the same slot overwrites it)."""
        return self._last_loss

    def grad_views(self):
        """(parameter, view of the averaged-gradient arena) pairs."""
        return list(zip(self.params, self.avg_views))

    def updated_params(self):
        return self.params

    def kernels_per_step(self):
        return self.kernel_count

    def outputs(self):
        """Logits of the most recent step (static buffer: the next step overwrites it)."""
        return self.graphs[self.last_key]["buf"]["logits"]

    def logits(self, N):
        """Logits of the most recent step with batch size N (static buffer)."""
        return self.graphs[(N, self.last_key[1])]["buf"]["logits"]
