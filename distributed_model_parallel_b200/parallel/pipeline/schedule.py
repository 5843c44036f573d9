"""Micro-batch pipeline execution: naive (one batch in flight = the reference), GPipe, and 1F1B.

The reference's "scheduler" is three hand-written loops (train_header / train_medium / train_last,
/root/reference/code/distributed_training/utils.py:34-210) that keep exactly one mini-batch in flight, so
GPU utilisation is bounded by 1/W and it measures 4x slower than DataParallel (Readme.md:283-286).
``PipelineStage`` runs any stage of a W-stage pipeline from a generated step order
(csrc/runtime/pipeline.cpp::dmp_schedule_generate): activations cross stage boundaries over NCCL/gloo p2p,
sends and receives that meet between two stages are batched, receives are posted before the compute they
feed where the schedule allows, and every stage owns a private optimizer like the reference's stages do.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, Dict, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from ... import _native
from ...utils.tracing import nvtx_range
from . import p2p
from .partition import partition_balanced  # noqa: F401  (registers the runtime signatures)

SCHEDULES = {"naive": 0, "gpipe": 1, "1f1b": 2}
_DEBUG = bool(int(__import__("os").environ.get("DMP_PP_DEBUG", "0")))


def generate_schedule(kind: str, num_stages: int, stage: int, num_microbatches: int) -> List[int]:
    """Step order of one stage: +(m+1) = forward of micro-batch m, -(m+1) = its backward."""
    out = (C.c_int32 * (2 * num_microbatches))()
    n = _native.runtime().dmp_schedule_generate(SCHEDULES[kind], num_stages, stage, num_microbatches, out, 2 * num_microbatches)
    if n < 0:
        raise ValueError("schedule %r is invalid for W=%d stage=%d M=%d (code %d)" % (kind, num_stages, stage, num_microbatches, n))
    return [int(out[i]) for i in range(n)]


def max_in_flight(order: List[int]) -> int:
    """Largest number of micro-batches whose forward has run but whose backward has not, over one stage's step order.
    Micro-batches retire in FIFO order in every schedule generated here, so this is also the number of activation slots
    (and captured CUDA-graph pairs) the stage needs: M for GPipe, min(M, W - stage) for 1F1B."""
    live = peak = 0
    for op in order:
        live += 1 if op > 0 else -1
        peak = max(peak, live)
    return peak


def simulate_schedule(kind: str, num_stages: int, num_microbatches: int, tf: float = 1.0, tb: float = 2.0):
    """(makespan, peak in-flight activations per stage) from the native event simulator; raises on deadlock."""
    peak = (C.c_int32 * num_stages)()
    t = _native.runtime().dmp_schedule_simulate(SCHEDULES[kind], num_stages, num_microbatches, tf, tb, peak)
    if t < 0:
        raise RuntimeError("schedule %r deadlocks or is invalid (code %s)" % (kind, t))
    return t, [int(x) for x in peak]


class PipelineStage:
    """One stage of a layer-split pipeline.

    ``loss_on`` selects where the loss is computed: ``"last"`` (targets travel to the last stage) or ``"first"``
    (logits are shipped back to stage 0 and dlogits return, the reference's ring, utils.py:56-58)."""

    def __init__(self, module: nn.Module, stage: int, num_stages: int, group=None, optimizer=None,
                 loss_fn: Optional[Callable] = None, schedule: str = "1f1b", loss_on: str = "last", device=None,
                 use_graphs: Optional[bool] = None):
        self.module, self.stage, self.W = module, stage, num_stages
        self.group = group
        self.optimizer = optimizer
        self.loss_fn = loss_fn
        self.schedule = schedule
        self.loss_on = loss_on
        self.first, self.last = stage == 0, stage == num_stages - 1
        if loss_on not in ("first", "last"):
            raise ValueError("loss_on must be 'first' or 'last'")
        if loss_on == "first" and schedule == "1f1b" and num_stages > 1:
            # with the loss on stage 0 every warm-up forward of stage 0 waits for logits from the last stage, while
            # 1F1B has the middle stages return gradients before stage 0 is ready to take them: use gpipe or naive
            raise ValueError("schedule '1f1b' needs loss_on='last' (the logits ring of the reference only fits naive/gpipe)")
        self.device = torch.device(device) if device is not None else next(module.parameters()).device
        ranks = dist.get_process_group_ranks(group) if group is not None else list(range(dist.get_world_size()))
        self.prev = ranks[stage - 1] if not self.first else None
        self.next = ranks[stage + 1] if not self.last else None
        self.rank0, self.rank_last = ranks[0], ranks[-1]
        self._in_meta: Optional[p2p.TensorMeta] = None   # per-sample-batch meta of the stage input (learned from a header)
        self._out_meta: Optional[p2p.TensorMeta] = None
        self._logits_meta: Optional[p2p.TensorMeta] = None
        self._probe_key = None
        self._keepalive, self._send_works = [], []
        # CUDA-graph every micro-batch's stage forward and backward (static shapes): a micro-batch step is then two graph
        # launches instead of hundreds of eager kernel launches, which is what makes GPipe/1F1B pay off on small models
        if use_graphs is None:
            use_graphs = os.environ.get("DMP_PIPELINE_GRAPHS", "1") != "0"
        self.use_graphs = bool(use_graphs) and self.device.type == "cuda" and loss_on == "last"
        self._slots, self._slots_key, self._gstream = None, None, None
        self._slot_cache: Dict[tuple, list] = {}   # captured micro-batch graphs per shape key (full batch, short last batch)
        self.last_train_hits = None                # device scalar: top-1 hits of the last train_step (loss-owning stage)
        self.last_eval_targets = None              # targets of the last eval_step on the loss-owning stage
        self.stats = {"steps": 0, "p2p_messages": 0, "graph_replays": 0, "graph_builds": 0}
        if self.device.type == "cuda" and num_stages > 1:
            self._warmup_p2p()

    def _warmup_p2p(self):
        """Establish the NCCL p2p connections of every neighbour pair with a SYMMETRIC grouped send+recv on both sides.
        NCCL sets up a pair's send and receive connectors on first use and the two ranks must agree on what is being
        connected; the schedules later issue asymmetric groups (one side send+recv, the other only recv), which fails
        on a cold pair ("Message truncated" inside ncclGroupEnd)."""
        def pair(peer):
            a = torch.zeros(8, device=self.device)
            b = torch.empty(8, device=self.device)
            p2p.wait_all(p2p.batched([("send", a, peer), ("recv", b, peer)], self.group))

        for parity in (0, 1):                      # even pairs first, then odd pairs: no rank is in two exchanges at once
            if self.stage % 2 == parity and self.next is not None:
                pair(self.next)
            if self.stage % 2 != parity and self.prev is not None:
                pair(self.prev)
        if self.W > 2:                             # the logits ring / target hand-off pair (first <-> last)
            if self.first:
                pair(self.rank_last)
            elif self.last:
                pair(self.rank0)
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------------------------------ control plane
    CTRL_LEN = 4 + p2p.HEADER_LEN

    def _probe_out_meta(self, in_meta: "p2p.TensorMeta") -> "p2p.TensorMeta":
        """Shape/dtype/layout of this stage's output for inputs like ``in_meta`` (one forward on zeros, eval mode,
        no grad -- only when the input meta is new).  Keeps every data-path transfer header-free."""
        key = (in_meta.shape[1:], in_meta.dtype, in_meta.channels_last)
        if self._probe_key == key and self._out_meta is not None:
            return self._out_meta.with_batch(in_meta.shape[0])
        was_training = self.module.training
        self.module.eval()
        with torch.no_grad():
            x = in_meta.with_batch(min(2, in_meta.shape[0])).empty(self.device).zero_()
            y = self.module(x)
        self.module.train(was_training)
        self._probe_key = key
        self._out_meta = p2p.TensorMeta.of(y)
        return self._out_meta.with_batch(in_meta.shape[0])

    def _control(self, info: Optional[List[int]], first_input: Optional[torch.Tensor]):
        """Per-step control message relayed stage to stage before any data moves:
        {batch, microbatches, training, 0, meta of the receiving stage's input}.  Returns (batch, M, training)."""
        if self.first:
            in_meta = p2p.TensorMeta.of(first_input)
            B, M, tr = info
        else:
            t = torch.empty(self.CTRL_LEN, dtype=torch.int64, device=self.device)
            dist.recv(t, self.prev, group=self.group)
            v = t.cpu()
            B, M, tr = int(v[0]), int(v[1]), int(v[2])
            in_meta = p2p.TensorMeta.unpack(v[4:])
        self._in_meta = in_meta
        out_meta = self._probe_out_meta(in_meta)
        if not self.last:
            msg = torch.cat([torch.tensor([B, M, tr, 0], dtype=torch.int64), out_meta.pack()]).to(self.device)
            dist.send(msg, self.next, group=self.group)
        if self.loss_on == "first" and self.W > 1:  # stage 0 must know the logits' meta to post its receive
            if self.last:
                dist.send(out_meta.pack().to(self.device), self.rank0, group=self.group)
            elif self.first:
                h = torch.empty(p2p.HEADER_LEN, dtype=torch.int64, device=self.device)
                dist.recv(h, self.rank_last, group=self.group)
                self._logits_meta = p2p.TensorMeta.unpack(h.cpu())
        return B, M, tr

    @staticmethod
    def _split_sizes(batch: int, m: int) -> List[int]:
        base, rem = divmod(batch, m)
        return [base + (1 if i < rem else 0) for i in range(m) if base + (1 if i < rem else 0) > 0]

    # ------------------------------------------------------------------------------------------ data plane
    def _exchange(self, send, recv_meta, recv_peer, into: Optional[torch.Tensor] = None):
        """One grouped p2p round: optionally send ``send = (tensor, peer)`` and receive a tensor of ``recv_meta`` from
        ``recv_peer``.  Grouping (batch_isend_irecv) lets opposite-direction transfers between two stages proceed
        together; the received tensor is returned once it has landed, the send is reaped at the end of the step."""
        ops, buf = [], None
        if send is not None:
            wire = p2p.wire_view(send[0].detach())
            self._keepalive.append(wire)
            ops.append(("send", wire, send[1]))
        if recv_meta is not None:
            buf = into if into is not None else recv_meta.empty(self.device)  # `into`: a static CUDA-graph input buffer
            ops.append(("recv", p2p.wire_view(buf), recv_peer))
        if _DEBUG:
            print("[pp stage %d] exchange %s" % (self.stage, [(k, tuple(t.shape), str(t.dtype), t.is_contiguous(), peer) for k, t, peer in ops]), flush=True)
        works = p2p.batched(ops, self.group)
        self.stats["p2p_messages"] += len(ops)
        if recv_meta is not None and len(works) == len(ops):
            works[-1].wait()                      # one work per op (gloo): only the receive gates the compute
            self._send_works.extend(works[:-1])   # sends are reaped at the end of the step
        elif recv_meta is not None:
            for w in works:                       # a coalesced NCCL group completes as a unit
                w.wait()
        else:
            self._send_works.extend(works)
        return buf

    # ------------------------------------------------------------------------------------------ graphed micro-batch slots
    def _zero_param_grads(self):
        for p in self.module.parameters():
            if p.grad is not None:
                p.grad.zero_()

    def _build_slots(self, mb: int, M: int, B: int, n_slots: int):
        """Captures, for each of the ``n_slots`` micro-batches that can be in flight on this stage at once (M for GPipe,
        at most W - stage for 1F1B), one CUDA graph of this stage's forward (plus the loss on the last stage) and one of its
        backward, on static input / target / incoming-gradient buffers.  Micro-batch m uses slot m % n_slots."""
        dev = self.device
        in_meta = self._in_meta.with_batch(mb)
        scale = float(mb) / float(B)
        # ONE stream for the warm-up and every capture: autograd runs a parameter's AccumulateGrad node on the stream that was
        # current when the node was created (the warm-up) and, at the end of backward, joins that stream into the caller's.
        # With gradients delivered straight into the optimizer arena those nodes receive nothing, so a different stream would
        # never have entered the capture and the join would be an illegal dependency on uncaptured work.
        if self._gstream is None:
            self._gstream = torch.cuda.Stream(device=dev)
        side = self._gstream
        side.wait_stream(torch.cuda.current_stream(dev))

        def make_x():
            x = in_meta.empty(dev)
            x.zero_()
            if not self.first and x.is_floating_point():
                x.requires_grad_(True)
            return x

        # The warm-up passes run in train mode on zeros: they must not leak into BatchNorm running statistics / counters.
        saved_buffers = [b.detach().clone() for b in self.module.buffers()]
        with torch.cuda.stream(side):  # eager warm-up so that lazy initialisation does not end up inside a capture
            for _ in range(2):
                x = make_x()
                y = self.module(x)
                if self.last:
                    y = self.loss_fn(y, torch.zeros(mb, dtype=torch.int64, device=dev)) * scale
                    y.backward()
                else:
                    torch.autograd.backward(y, torch.zeros_like(y))
            self._zero_param_grads()
            with torch.no_grad():
                for b, v in zip(self.module.buffers(), saved_buffers):
                    b.copy_(v)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        # One private memory pool per slot.  A backward capture returns the slot's saved activations to its pool, so with a
        # shared pool the temporaries of another slot's graphs would be laid over them -- and under 1F1B a slot is re-filled
        # (forward of micro-batch m + n_slots) while its neighbour's backward is still to run.  Inside a slot the forward is
        # captured before the backward, and they replay in that order, so sharing there is safe.
        slots = []
        for m in range(n_slots):
            pool = torch.cuda.graph_pool_handle()
            sl = {"x": make_x(), "t": torch.zeros(mb, dtype=torch.int64, device=dev) if self.last else None}
            sl["gf"] = torch.cuda.CUDAGraph()
            with torch.cuda.graph(sl["gf"], pool=pool, stream=side):
                y = self.module(sl["x"])
                if self.last:
                    with torch.no_grad():
                        sl["hits"] = (y.detach().argmax(1) == sl["t"]).sum()
                    y = self.loss_fn(y, sl["t"]) * scale
            sl["y"] = y
            sl["g"] = None if self.last else p2p.TensorMeta.of(y).empty(dev).zero_()
            sl["gb"] = torch.cuda.CUDAGraph()
            with torch.cuda.graph(sl["gb"], pool=pool, stream=side):
                if self.last:
                    y.backward()
                else:
                    torch.autograd.backward(y, sl["g"])
            sl["dx"] = sl["x"].grad if not self.first else None
            slots.append(sl)
        self._zero_param_grads()
        self._slots, self._slots_key = slots, self._graph_key(mb, n_slots, B)
        self.stats["graph_builds"] += 1
        if len(self._slot_cache) >= 3:  # full batch + short last batch is the expected population; bound the pools held
            self._slot_cache.pop(next(iter(self._slot_cache)))
        self._slot_cache[self._slots_key] = slots

    def _graph_key(self, mb: int, n_slots: int, B: int):
        """Everything a captured micro-batch graph has baked in: shapes, the loss scale mb/B, and the addresses of the gradient
        tensors it accumulates into (an optimizer that re-homes .grad, or zero_grad(set_to_none=True), forces a re-capture)."""
        meta = self._in_meta.with_batch(mb)
        grads = tuple(p.grad.data_ptr() if p.grad is not None else 0 for p in self.module.parameters())
        return (mb, n_slots, B, tuple(meta.shape), meta.dtype, self.module.training, hash(grads))

    # ------------------------------------------------------------------------------------------ one training step
    def train_step(self, inputs: Optional[torch.Tensor] = None, targets: Optional[torch.Tensor] = None,
                   num_microbatches: int = 1):
        """Runs forward+backward of every micro-batch in schedule order, then the stage's optimizer step.
        Stage 0 passes ``inputs`` and ``targets`` (targets are forwarded to where the loss lives).  Returns the mean
        loss on the stage that computes it, else None."""
        self.module.train()
        if self.first:
            inputs = inputs.to(self.device)
            B, M, _ = self._control([int(inputs.shape[0]), int(num_microbatches), 1], inputs)
        else:
            B, M, _ = self._control(None, None)
        sizes = self._split_sizes(B, M)
        M = len(sizes)
        self.last_batch = B
        kind = self.schedule if M > 1 else "naive"
        order = generate_schedule(kind, self.W, self.stage, M)
        loss_here = (self.loss_on == "last" and self.last) or (self.loss_on == "first" and self.first)
        self._keepalive, self._send_works = [], []

        tgt_chunks = None
        if self.loss_on == "last" and self.W > 1:
            if self.first:
                dist.send(targets.to(self.device).contiguous(), self.rank_last, group=self.group)
            elif self.last:
                t = torch.empty(B, dtype=torch.int64, device=self.device)
                dist.recv(t, self.rank0, group=self.group)
                tgt_chunks = list(t.split(sizes))
        elif loss_here and targets is not None:
            tgt_chunks = list(targets.to(self.device).split(sizes))
        in_chunks = list(inputs.split(sizes)) if self.first else None
        out_meta = self._out_meta

        saved_in: Dict[int, torch.Tensor] = {}
        saved_out: Dict[int, torch.Tensor] = {}
        total_loss = None
        hits = None
        pending = None  # (tensor, peer) produced by the previous step, sent together with the next receive
        graphed = self.use_graphs and len(set(sizes)) == 1
        if graphed:
            n_slots = max_in_flight(order)
            key = self._graph_key(sizes[0], n_slots, B)
            if self._slots_key != key and key in self._slot_cache:
                self._slots, self._slots_key = self._slot_cache[key], key
            elif self._slots_key != key:
                try:
                    self._build_slots(sizes[0], M, B, n_slots)
                except Exception as e:  # e.g. the module cannot be captured: keep training, eagerly (every stage decides alone;
                    import warnings     # graphs change nothing on the wire)
                    warnings.warn("pipeline stage %d: CUDA-graph capture failed (%s); running micro-batches eagerly" % (self.stage, e))
                    torch.cuda.synchronize(self.device)
                    self._slots, self._slots_key, self.use_graphs, graphed = None, None, False, False
                    self._zero_param_grads()
        for op in order:
            m = abs(op) - 1
            if graphed:  # ---------------------------------------------------------------- two graph launches per micro-batch
                sl = self._slots[m % n_slots]
                if op > 0:
                    if m >= n_slots:
                        p2p.wait_all(self._send_works)  # the slot's previous occupant: its sends must have left the buffers
                        self._send_works = []
                    if self.first:
                        if pending is not None:
                            self._exchange(pending, None, None)
                            pending = None
                        sl["x"].detach().copy_(in_chunks[m])
                    else:
                        self._exchange(pending, self._in_meta.with_batch(sizes[m]), self.prev, into=sl["x"].detach())
                        pending = None
                    if self.last:
                        sl["t"].copy_(tgt_chunks[m])
                    with nvtx_range("pp.s%d.F%d" % (self.stage, m)):
                        sl["gf"].replay()
                    if self.last:
                        total_loss = sl["y"].detach().clone() if total_loss is None else total_loss + sl["y"].detach()
                        hits = sl["hits"].clone() if hits is None else hits + sl["hits"]
                    else:
                        pending = (sl["y"], self.next)
                else:
                    if self.last:
                        if pending is not None:
                            self._exchange(pending, None, None)
                            pending = None
                    else:
                        self._exchange(pending, p2p.TensorMeta.of(sl["g"]), self.next, into=sl["g"])
                        pending = None
                    with nvtx_range("pp.s%d.B%d" % (self.stage, m)):
                        sl["gb"].replay()
                    if not self.first:
                        pending = (sl["dx"], self.prev)
                self.stats["graph_replays"] += 1
                continue
            if op > 0:  # ---------------------------------------------------------------- forward of micro-batch m
                if self.first:
                    if pending is not None:
                        self._exchange(pending, None, None)
                        pending = None
                    x = in_chunks[m]
                else:
                    x = self._exchange(pending, self._in_meta.with_batch(sizes[m]), self.prev)
                    pending = None
                    if x.is_floating_point():
                        x.requires_grad_(True)
                with nvtx_range("pp.s%d.F%d" % (self.stage, m)):
                    y = self.module(x)
                saved_in[m] = x
                if not self.last:
                    saved_out[m] = y
                    pending = (y, self.next)
                elif self.loss_on == "last":
                    loss = self.loss_fn(y, tgt_chunks[m]) * (sizes[m] / B)
                    with torch.no_grad():
                        h = (y.detach().argmax(1) == tgt_chunks[m]).sum()
                        hits = h if hits is None else hits + h
                    saved_out[m] = loss
                    total_loss = loss.detach() if total_loss is None else total_loss + loss.detach()
                else:  # ring: logits go back to stage 0, which owns targets and loss
                    saved_out[m] = y
                    pending = (y, self.rank0)
                if self.first and self.loss_on == "first" and self.W > 1:
                    # stage 0 turns the logits of micro-batch m into dlogits as soon as they arrive
                    lg = self._exchange(pending, self._logits_meta.with_batch(sizes[m]), self.rank_last)
                    pending = None
                    lg.requires_grad_(True)
                    loss = self.loss_fn(lg, tgt_chunks[m]) * (sizes[m] / B)
                    total_loss = loss.detach() if total_loss is None else total_loss + loss.detach()
                    with torch.no_grad():
                        h = (lg.detach().argmax(1) == tgt_chunks[m]).sum()
                        hits = h if hits is None else hits + h
                    loss.backward()
                    self._exchange((lg.grad, self.rank_last), None, None)
            else:  # ---------------------------------------------------------------------- backward of micro-batch m
                y = saved_out.pop(m)
                x = saved_in.pop(m)
                if self.last and self.loss_on == "last":
                    if pending is not None:
                        self._exchange(pending, None, None)
                        pending = None
                    y.backward()
                else:
                    src = self.rank0 if self.last else self.next
                    g = self._exchange(pending, p2p.TensorMeta.of(y), src)
                    pending = None
                    torch.autograd.backward(y, g)
                if not self.first:
                    pending = (x.grad, self.prev)
        if pending is not None:
            self._exchange(pending, None, None)
        p2p.wait_all(self._send_works)
        self._keepalive, self._send_works = [], []
        if self.optimizer is not None:
            self.optimizer.step()
            if graphed:
                self.optimizer.zero_grad(set_to_none=False)  # captured graphs accumulate into these very tensors
            else:
                self.optimizer.zero_grad()
        self.stats["steps"] += 1
        self.last_train_hits = hits if loss_here else None
        return total_loss if loss_here else None

    # ------------------------------------------------------------------------------------------ evaluation
    @torch.no_grad()
    def eval_step(self, inputs: Optional[torch.Tensor] = None, targets: Optional[torch.Tensor] = None):
        """Forward-only pass of one batch (the reference's val_header / val_medium / val_last, utils.py:81-210).
        Returns (loss, logits) on the stage that owns the loss, else (None, None); that stage also keeps the batch's targets
        in ``last_eval_targets`` (with ``loss_on="last"`` they were shipped there), for accuracy bookkeeping."""
        self.module.eval()
        self.last_eval_targets = None
        if self.first:
            inputs = inputs.to(self.device)
            B, _, _ = self._control([int(inputs.shape[0]), 1, 0], inputs)
        else:
            B, _, _ = self._control(None, None)
        self._keepalive, self._send_works = [], []
        x = inputs if self.first else self._exchange(None, self._in_meta.with_batch(B), self.prev)
        y = self.module(x)
        result = (None, None)
        if not self.last:
            self._exchange((y, self.next), None, None)
        if self.loss_on == "last":
            if self.W > 1 and self.first:
                dist.send(targets.to(self.device).contiguous(), self.rank_last, group=self.group)
            if self.last:
                if self.W > 1:
                    t = torch.empty(B, dtype=torch.int64, device=self.device)
                    dist.recv(t, self.rank0, group=self.group)
                else:
                    t = targets.to(self.device)
                self.last_eval_targets = t
                result = (self.loss_fn(y, t), y)
        else:
            if self.last and self.W > 1:
                self._exchange((y, self.rank0), None, None)
            if self.first:
                lg = self._exchange(None, self._logits_meta.with_batch(B), self.rank_last) if self.W > 1 else y
                self.last_eval_targets = targets.to(self.device)
                result = (self.loss_fn(lg, self.last_eval_targets), lg)
        p2p.wait_all(self._send_works)
        self._keepalive, self._send_works = [], []
        return result
