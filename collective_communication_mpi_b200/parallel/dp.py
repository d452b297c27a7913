"""Data-parallel gradient reduction: all-reduce fused with the 1/dp scale (and an optional
cast), bucketed, and overlapped with the rest of backward on a side stream.

The reference only *defines* the data-parallel communicator (``dp_comm``, func_impl.py:60-62)
and checks it with one ``Allreduce`` (tests/test_get_info.py:39); the gradient all-reduce itself
is implied. Here it is a first-class path:

* every weight gradient lives in ONE flat bf16 buffer in the symmetric heap, so the dW GEMMs
  write their result where the collective reads it (no packing copies);
* a bucket is reduced as soon as its last gradient has been produced, by the all-reduce kernel
  of csrc/collectives.cu (NVLS ``multimem.ld_reduce``/``multimem.st`` with fp32 accumulation when
  the switch supports it, two-shot P2P otherwise), in place, with ``scale = 1/dp`` applied in the
  same pass;
* the kernel runs on its own stream next to the remaining backward GEMMs, on SMs of its own.

SM partition (why ``max_blocks`` and ``layer.bwd_num_sms`` exist). A reduction block (512 threads
x 82-128 registers) fills an SM's register file, so it never shares an SM with a GEMM CTA. The
fused GEMM+collective kernels of a model-parallel group assign tiles and push work statically to
their CTAs, which wait on peers: every CTA must be resident. If a rank's reduction took SMs the
next fused GEMM needed while its data-parallel partner had them the other way round, the two
kernels of the two groups would wait on each other for ever. Therefore, whenever buckets overlap a
model-parallel backward (mp > 1 and dp > 1): the reduction gets at most ``max_blocks`` blocks,
launched as 2-CTA clusters so that it takes whole SM pairs (the GEMM's CTA pairs need both SMs of a
pair), and the layers' backward fused kernels are sized to the remaining SMs. Together they always
fit, whatever order the hardware starts them in. The last bucket has nothing to hide behind: it is
launched on the main stream, after the kernels before it AND after the buckets still reducing on the side
stream (collectives of one communicator share its barrier words and must not overlap), and may use every SM.

Optimizer. With NVLS the reduction IS the optimizer step (``fused_update``): the weights live in a
second flat symmetric buffer with the gradients' layout, and the kernel of a bucket sums its slice of
every replica's gradients in the switch, updates its slice of the weights (stochastic rounding,
csrc/sr.cuh) and multicasts the new weights to all replicas -- no reduced gradient is written, no
optimizer pass runs, replicas stay bit-identical. Because that kernel rewrites a layer's weights
during backward, layers attached to such a sink compute dx (which reads W) before dW
(``layer.dw_last``). Without NVLS the gradients are all-reduced in place and ``apply`` runs the
stand-alone update kernel per layer.
"""
from __future__ import annotations

import os
from typing import Dict, List

import torch

from .. import _native as N
from ..ops import collectives as CO


class GradSink:
    def __init__(self, dp_comm, layers, *, bucket_mb: int = 64, overlap: bool = True, max_blocks: int = 16):
        self.comm = dp_comm
        self.dp = dp_comm.Get_size()
        self.layers = list(layers)
        params = [l.weight for l in self.layers]
        self.total = sum(p.numel() for p in params)
        pad = lambda n: (n + 127) // 128 * 128  # noqa: E731  keep every view 256-byte aligned
        self.on_gpu = params[0].is_cuda
        if self.on_gpu:
            self.flat = dp_comm.symm_empty(sum(pad(p.numel()) for p in params), torch.bfloat16)
            self.dev = dp_comm.device_comm()
            self.flat_off = self.dev.heap.offset_of(self.flat)
            self.use_nvls = self.dev.has_mc and self.dp > 2      # measured: the P2P two-shot wins between two GPUs
            self.tmp = dp_comm.symm_empty(self.flat.numel(), torch.bfloat16) if self.dp > 1 and not self.use_nvls else None
        else:
            # no GPU in this process (CPU test-suite): same buckets, reduced by the library
            # all-reduce of the shared-memory transport when they complete
            self.flat = torch.zeros(sum(pad(p.numel()) for p in params), dtype=torch.bfloat16)
            self.dev, self.flat_off, self.use_nvls, self.tmp = None, 0, False, None
        self._views: Dict[int, torch.Tensor] = {}
        self._bucket_of: Dict[int, int] = {}
        self.buckets: List[dict] = []
        # buckets follow the order gradients appear in backward: last layer first
        off, cur = 0, None
        limit = bucket_mb << 20
        for layer in reversed(self.layers):
            p = layer.weight
            n = pad(p.numel())
            if cur is None or (cur["hi"] - cur["lo"]) * 2 >= limit:
                cur = {"lo": off, "hi": off, "want": 0, "have": 0}
                self.buckets.append(cur)
            self._views[id(p)] = self.flat[off:off + p.numel()].view_as(p)
            self._bucket_of[id(p)] = len(self.buckets) - 1
            cur["hi"] = off + n
            cur["want"] += 1
            off += n
            layer.grad_sink = self
        self.overlap = overlap and self.dp > 1 and self.on_gpu
        self.stream = torch.cuda.Stream() if self.overlap else None
        self.max_blocks = max_blocks & ~1
        self.algo = N.AR_NVLS if self.use_nvls else N.AR_TWOSHOT
        self.launches = 0
        self.lr, self.step_no, self.seed = 0.0, 0, 0x5EED
        # the reduction doubles as the optimizer step when the switch can do it (see the module docstring)
        self.fused_update = bool(self.on_gpu and self.dp > 1 and self.dev.has_mc and
                                 os.environ.get("CCMPI_DP_FUSED_UPDATE", "1") != "0")
        self.wflat = None
        if self.fused_update:
            self.wflat = dp_comm.symm_empty(self.flat.numel(), torch.bfloat16)
            self.wflat.zero_()
            self.wflat_off = self.dev.heap.offset_of(self.wflat)
            for l in self.layers:       # the parameters move into the flat buffer, at their gradients' offsets
                v = self._views[id(l.weight)]
                off = int(v.storage_offset() - self.flat.storage_offset())
                wv = self.wflat[off:off + l.weight.numel()].view_as(l.weight)
                wv.copy_(l.weight.data)
                l.weight.data = wv
        # reserve the reduction's SMs in the backward fused kernels of a model-parallel group
        self.partitioned = self.overlap and any(getattr(l, "mp", 1) > 1 for l in self.layers)
        if self.partitioned:
            sms = self.comm._world.sm_count
            for l in self.layers:
                if hasattr(l, "bwd_num_sms"):
                    l.bwd_num_sms = (sms - self.max_blocks) & ~1
                else:       # layer kinds with one SM budget for all their kernels
                    l.num_sms = (sms - self.max_blocks) & ~1

    def view_for(self, param) -> torch.Tensor:
        return self._views[id(param)]

    def grads(self) -> List[torch.Tensor]:
        return [self._views[id(l.weight)] for l in self.layers]

    @property
    def updates_in_backward(self) -> bool:
        """True when a bucket's kernel rewrites weights while backward is still running (layers then
        read W for dx before they hand over dW)."""
        return self.fused_update

    def begin_step(self, lr: float, step_no: int):
        """Learning rate and step counter (stochastic-rounding stream) of the step that is about to run."""
        self.lr, self.step_no = float(lr), int(step_no)

    def ready(self, param):
        b = self.buckets[self._bucket_of[id(param)]]
        b["have"] += 1
        if b["have"] == b["want"]:
            b["have"] = 0
            self._reduce(b)

    def _reduce(self, b):
        if self.dp == 1:
            return
        # the bucket produced last (first layer's gradients) has nothing left to hide behind: give
        # it every SM; earlier buckets share the GPU with the remaining backward GEMMs
        last = b is self.buckets[-1]
        count = b["hi"] - b["lo"]
        if not self.on_gpu:
            view = self.flat[b["lo"]:b["hi"]]
            total = torch.empty_like(view)
            self.comm.Allreduce(view, total)
            view.copy_(total.float() * (1.0 / self.dp))
            self.launches += 1
            return
        off = self.flat_off + b["lo"] * 2
        tmp_off = self.dev.heap.offset_of(self.tmp) + b["lo"] * 2 if self.tmp is not None else 0
        cur = torch.cuda.current_stream()
        side = self.overlap and not last
        st = self.stream if side else cur
        if side:
            ev = torch.cuda.Event()
            ev.record(cur)
            st.wait_event(ev)
        elif self.overlap:
            # Two collectives on ONE communicator must never run at the same time: they share the communicator's
            # per-block barrier words (block b of every rank pairs up through pad[b]), so a bucket still reducing on
            # the side stream and this one would corrupt each other's epochs and hang (seen at dp = 4 after ~170
            # barriers). The last bucket therefore also waits for the side stream.
            cur.wait_stream(self.stream)
        blocks = 0 if not side else (self.max_blocks | N.BLOCKS_PAIRED if self.partitioned else self.max_blocks)
        if self.fused_update:
            CO.allreduce_sgd(self.dev.comms, off, self.wflat_off + b["lo"] * 2, count, self.lr / self.dp, base_index=b["lo"],
                             step=self.step_no, seed=self.seed, max_blocks=blocks, stream=st)
        else:
            CO.allreduce(self.dev.comms, N.BF16, N.SUM, off, off, count, tmp_off=tmp_off, algo=self.algo,
                         scale=1.0 / self.dp, max_blocks=blocks, stream=st)
        self.launches += 1

    def finish(self):
        """Make the reduced gradients visible to the current stream (call before the optimizer)."""
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self.stream)


    def apply(self):
        """The optimizer step for what the reductions left to do: nothing with ``fused_update``; otherwise one
        stand-alone update kernel per layer on the (already averaged) gradients, ``w = SR(w - lr * g)``."""
        self.finish()
        if self.fused_update:
            return
        from ..ops import elementwise as EW
        with torch.no_grad():
            for l in self.layers:
                g = self._views[id(l.weight)]
                base = int(g.storage_offset() - self.flat.storage_offset())     # position in the flat numbering
                EW.sgd_sr_(l.weight.data, g, self.lr, base_index=base, step=self.step_no, seed=self.seed)


def sgd_step(weights: List[torch.Tensor], grads: List[torch.Tensor], lr: float):
    """Plain round-to-nearest SGD (kept for comparisons; the training step uses ``GradSink.apply``)."""
    with torch.no_grad():
        torch._foreach_add_(weights, grads, alpha=-lr)
