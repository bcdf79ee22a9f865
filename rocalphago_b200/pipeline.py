"""Two-stage evaluation pipeline: features of batch k+1 overlap the convolutions of batch k.

`rag_eval_batch` runs the two halves of the hot path back to back on one stream.  They stress
different parts of the SM -- the feature kernels are integer/latency work (a few long ladder
searches leave most of the chip idle), the convolutions are tensor-core work -- so a service
that evaluates a stream of batches runs them on two CUDA streams with double-buffered
intermediates.  This is host-side plumbing over the same C-ABI entry points
(`rag_features` with RAG_LAYOUT_CONV_IN, then `rag_net_forward_rows`); results are identical.
"""
import ctypes

import numpy as np
import torch

from . import _cabi
from .preprocessing import LAYOUT_CONV_IN
from .state import STATE_WORDS, _ptr, _require_cuda


class _Slot(object):
    def __init__(self, L, handle, n, size, device, out_cols, value_handle=None):
        rows = int(L.rag_net_input_rows(handle, n))
        ch = int(L.rag_net_input_channels(handle))
        self.packed = torch.empty((n, STATE_WORDS), dtype=torch.int32, device=device)
        self.rows = torch.empty((rows, ch), dtype=torch.float16, device=device)        # 16-bit payload, either dtype
        self.fws = torch.empty(int(L.rag_features_workspace_bytes(n)), dtype=torch.uint8, device=device)
        self.out = torch.empty((n, out_cols) if out_cols else (n,), dtype=torch.float32, device=device)
        self.host_out = torch.empty_like(self.out, device="cpu").pin_memory()
        self.vrows = self.vout = self.host_vout = None
        if value_handle is not None:
            self.vrows = torch.empty((int(L.rag_net_input_rows(value_handle, n)), int(L.rag_net_input_channels(value_handle))),
                                     dtype=torch.float16, device=device)
            self.vout = torch.empty((n,), dtype=torch.float32, device=device)
            self.host_vout = torch.empty_like(self.vout, device="cpu").pin_memory()
        self.stream = torch.cuda.Stream(device=device)      # this slot's feature stage runs here
        self.feat_done = torch.cuda.Event()
        self.net_done = torch.cuda.Event()
        self.out_copied = torch.cuda.Event()
        self.used = False
        self.uses = 0
        self.graphs = {}


class EvalPipeline(object):
    """Evaluates fixed-size batches of packed states with a policy or value network.

        pipe = EvalPipeline(net, batch_size)
        for packed in batches:                 # device or pinned-host int32 (n, 192) tensors
            ticket = pipe.submit(packed, to_host=True)
        out = pipe.result(ticket)              # numpy / tensor of the ticket's batch

    `slots` batches are in flight at most; a result stays valid until `slots` further batches have
    been submitted.
    """

    def __init__(self, net, batch_size, device=None, size=None, value_net=None, graphs=False, slots=None):
        """`value_net` (optional, a CNNValue): evaluated on the same positions in the same pass --
        the analysis and the ladder search are shared, each network gets its own planes.
        `graphs`: replay each stage as a CUDA graph from the second use of a slot on; cuts the host's
        cost of a step from ~450 us to ~60 us.  (Per-layer profiling -- rag_net_set_profiling -- must
        be off then.)
        `slots`: batches in flight (default 3 for large batches -- two stall the forwards whenever a batch's
        ladder stage outlasts one forward -- and 8 below 2,048 positions).  Each slot
        has its own stream for the feature stage, so with more slots several ladder searches --
        latency-bound, a few warps each at small batch sizes -- run side by side under the forwards;
        forwards run in submission order, so the slots must cover the longest search (~2.5 ms against
        a 0.7 ms forward at 256 positions: 4 slots give 236k positions/s there, 8 give 331k)."""
        _require_cuda()
        self.net, self.model, self.n = net, net.model, int(batch_size)
        self.value_net = value_net
        self.graphs = bool(graphs)
        self.device = torch.device(device or "cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.size = int(size or self.model.board)
        self.L = _cabi.lib()
        self.handle = self.model._ensure_handle(self.device)
        self.plane_src = np.ascontiguousarray(net.preprocessor.plane_src, np.int32)
        if len(self.plane_src) != self.model.input_dim:
            raise ValueError("feature list gives %d planes, the network takes %d" % (len(self.plane_src), self.model.input_dim))
        cols = self.model.board ** 2 if self.model.head == "policy" else 0
        self.vhandle = self.vsrc = None
        if value_net is not None:
            if value_net.model.board != self.model.board:
                raise ValueError("policy and value networks must have the same board size")
            self.vhandle = value_net.model._ensure_handle(self.device)
            self.vsrc = np.ascontiguousarray(value_net.preprocessor.plane_src, np.int32)
            self.vchannels = int(self.L.rag_net_input_channels(self.vhandle))
            self.vdtype_code = 0 if value_net.model.dtype == "bf16" else 1
        self.trace = None       # set to [] to collect, per batch, timing events [start, analysis end, features end, forward start, forward end] (tools/pipeline_timeline.py)
        self.nslots = int(slots) if slots else (3 if self.n >= 2048 else 8)
        if self.nslots < 2:
            raise ValueError("at least two slots")
        with torch.cuda.device(self.device):
            self.slots = [_Slot(self.L, self.handle, self.n, self.size, self.device, cols, self.vhandle)
                          for _ in range(self.nslots)]
            nws = int(self.L.rag_net_workspace_bytes(self.handle, self.n))
            if self.vhandle is not None:
                nws = max(nws, int(self.L.rag_net_workspace_bytes(self.vhandle, self.n)))
            self.nws = torch.empty(nws, dtype=torch.uint8, device=self.device)
            # All streams have the same priority.  (Giving the convolutions a higher one looked
            # attractive, but with the CTA-pair convolution kernel -- a cluster launch -- and the
            # persistent ladder kernel in flight together, prioritised streams stopped making
            # progress on B200 / driver 580; equal priorities never did in any run.)
            self.s_net, self.s_copy, self.s_stat = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
        self.channels = int(self.L.rag_net_input_channels(self.handle))
        self.dtype_code = 0 if self.model.dtype == "bf16" else 1
        self.count = 0

    # ---- the three device stages of one batch (all pointers are the slot's own: capturable) --------
    def _stage_analysis(self, slot, stream):
        _cabi.check(self.L.rag_features_analyze(_ptr(slot.packed), self.n, self.size, 2, _ptr(slot.fws), slot.fws.numel(),
                                                ctypes.c_void_p(stream.cuda_stream)))

    def _stage_ladders_emit(self, slot, stream):
        L, vp = self.L, ctypes.c_void_p
        st = vp(stream.cuda_stream)
        _cabi.check(L.rag_features_ladders(_ptr(slot.packed), self.n, self.size, _ptr(slot.fws), slot.fws.numel(), st))
        _cabi.check(L.rag_features_emit(_ptr(slot.fws), self.n, self.size, self.plane_src.ctypes.data_as(vp), len(self.plane_src),
                                        LAYOUT_CONV_IN, self.dtype_code, self.channels, _ptr(slot.rows), st))
        if self.vhandle is not None:
            _cabi.check(L.rag_features_emit(_ptr(slot.fws), self.n, self.size, self.vsrc.ctypes.data_as(vp), len(self.vsrc),
                                            LAYOUT_CONV_IN, self.vdtype_code, self.vchannels, _ptr(slot.vrows), st))

    def _stage_forward(self, slot, stream):
        L, st = self.L, ctypes.c_void_p(stream.cuda_stream)
        _cabi.check(L.rag_net_forward_rows(self.handle, _ptr(slot.rows), self.n, _ptr(slot.out), _ptr(self.nws), self.nws.numel(), st))
        if self.vhandle is not None:
            _cabi.check(L.rag_net_forward_rows(self.vhandle, _ptr(slot.vrows), self.n, _ptr(slot.vout), _ptr(self.nws),
                                               self.nws.numel(), st))

    def _run(self, slot, which, fn, stream):
        """Runs a stage on `stream` (already current): eagerly, or -- with graphs=True, from the second use
        of the slot on -- as a replay of its CUDA graph (one launch instead of up to fourteen)."""
        if not self.graphs or slot.uses < 1:
            fn(self, slot, stream)
            return
        g = slot.graphs.get(which)
        if g is None:
            g = torch.cuda.CUDAGraph()
            stream.synchronize()
            with torch.cuda.graph(g, stream=stream):
                fn(self, slot, torch.cuda.current_stream())
            slot.graphs[which] = g
        g.replay()

    @staticmethod
    def _trace_point(stream):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(stream)
        return [ev]

    def submit(self, packed, to_host=False):
        """Enqueue one batch; returns a ticket for `result`."""
        if tuple(packed.shape) != (self.n, STATE_WORDS):
            raise ValueError("expected a (%d, %d) int32 tensor" % (self.n, STATE_WORDS))
        k = self.count
        slot = self.slots[k % self.nslots]
        with torch.cuda.device(self.device):
            # stage 1 on the slot's own stream: copy the states in, analysis, ladder reader, emit pixel rows
            if packed.device.type != "cpu":
                slot.stream.wait_stream(torch.cuda.current_stream())          # whoever produced `packed`
            with torch.cuda.stream(slot.stream):
                if slot.used:
                    slot.stream.wait_event(slot.net_done)          # the forward that read this slot's buffers has finished
                slot.packed.copy_(packed, non_blocking=True)
                tr = self._trace_point(slot.stream) if self.trace is not None else None
                self._run(slot, "ana", EvalPipeline._stage_analysis, slot.stream)
                if tr: tr.append(self._trace_point(slot.stream)[0])
                self._run(slot, "feat", EvalPipeline._stage_ladders_emit, slot.stream)
                if tr: tr.append(self._trace_point(slot.stream)[0])
                slot.feat_done.record(slot.stream)
            # stage 2: convolutions + head
            with torch.cuda.stream(self.s_net):
                self.s_net.wait_event(slot.feat_done)
                if slot.used:
                    self.s_net.wait_event(slot.out_copied)         # the previous result of this slot has been copied out
                if tr: tr.append(self._trace_point(self.s_net)[0])
                self._run(slot, "net", EvalPipeline._stage_forward, self.s_net)
                if tr:
                    tr.append(self._trace_point(self.s_net)[0])
                    self.trace.append(tr)
                slot.net_done.record(self.s_net)
            self.s_copy.wait_event(slot.net_done)
            if to_host:
                with torch.cuda.stream(self.s_copy):
                    slot.host_out.copy_(slot.out, non_blocking=True)
                    if self.vhandle is not None:
                        slot.host_vout.copy_(slot.vout, non_blocking=True)
            slot.out_copied.record(self.s_copy)
        slot.used = True
        slot.uses += 1
        slot.to_host = to_host
        self.count += 1
        return k

    def result(self, ticket):
        """Waits for batch `ticket` (must be one of the last `slots` submitted) and returns its output:
        the pinned host tensor if it was submitted with to_host=True, else the device tensor."""
        if ticket < self.count - self.nslots or ticket >= self.count:
            raise ValueError("ticket %d is no longer (or not yet) in flight" % ticket)
        slot = self.slots[ticket % self.nslots]
        slot.out_copied.synchronize()
        with torch.cuda.device(self.device):       # did this batch's ladder kernel give up (watchdog)?  4 bytes on a stream of its own
            _cabi.check(self.L.rag_features_status(_ptr(slot.fws), ctypes.c_void_p(self.s_stat.cuda_stream)))
        out = slot.host_out if slot.to_host else slot.out
        if self.vhandle is not None:
            return out, (slot.host_vout if slot.to_host else slot.vout)
        return out

    def mark(self):
        """A CUDA event recorded behind everything submitted so far (for timing)."""
        ev = torch.cuda.Event(enable_timing=True)
        with torch.cuda.device(self.device):
            for st in [sl.stream for sl in self.slots] + [self.s_net]:
                self.s_copy.wait_stream(st)
            ev.record(self.s_copy)
            for sl in self.slots:
                sl.stream.wait_event(ev)        # later submissions start after the mark
        return ev

    def drain(self):
        for s in [sl.stream for sl in self.slots] + [self.s_net, self.s_copy]:
            s.synchronize()
