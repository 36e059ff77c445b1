"""Mirror of ``utils/train_eval_utils.py`` for the hot path: the training-step body
(:28-52) and the evaluation count (:83), on the fused kernels.

``FusedTrainer.step`` is one optimizer step: forward -> fused MSE(sum)+count+dL/det ->
backward into a flat fp32 gradient arena -> bucketed NCCL all-reduce overlapped with the rest
of backward (replacing DistributedDataParallel, train.py:121-122) -> one fused SGD-momentum
kernel over the flat parameter arena (torch.optim.SGD semantics, train.py:63,125-126).
Nothing in the step reads a value back to the host: the loss stays on the device.

``train_one_epoch`` / ``evaluate`` keep the reference signatures and return values.
"""
from __future__ import annotations

import os
import sys

import torch
import torch.distributed as dist

from .. import engine, ops
from .graph_cache import GraphCache
from .distributed_utils import get_world_size, is_dist_avail_and_initialized, is_main_process, reduce_value
from .dp import GradBuckets
from . import nvtx


class FusedTrainer:
    """Flat-arena SGD(momentum) trainer for a ``can_b200.model.CANNet``.

    DDP semantics are preserved: the loss is a SUM over the local batch, gradients are AVERAGED
    over ranks (train.py:121), the caller multiplies lr by world_size (train.py:25)."""

    def __init__(self, model, lr, momentum=0.95, process_group=None):
        self.model = model
        self.lr = float(lr)
        self.momentum = float(momentum)
        self.pg = process_group
        self.world = get_world_size()
        params = list(model.parameters())
        dev = params[0].device
        # Arena layout (the state_dict is untouched; only where each parameter's storage lives is chosen here):
        #   [ header: 64 floats, slot 0 = this rank's loss ][ frontend, parameter order ][ backend + head ]
        #   [ conv{1,2,3,6}_1 ][ conv{1,2,3,6}_2 ]
        # every parameter starts on a 256-byte boundary, so the kernels' 16-byte stores can write gradients in place and
        # the four conv{S}_1 / conv{S}_2 gradients are each one [4,512,512] block (engine._stacked_view).  The header
        # rides in the last gradient bucket: the loss needs no all-reduce of its own (distributed_utils.py:60-70).
        HDR, ALIGN = 64, 64
        order = list(range(engine.IDX_CTX)) + [engine.IDX_CTX + 2 * j for j in range(4)] + \
                [engine.IDX_CTX + 2 * j + 1 for j in range(4)]
        offs, off = {}, HDR
        for i in order:
            offs[i] = off
            off += (params[i].numel() + ALIGN - 1) // ALIGN * ALIGN
        total = off
        self.arena = torch.zeros(total, device=dev, dtype=torch.float32)
        self.garena = torch.zeros(total, device=dev, dtype=torch.float32)
        self.buf = torch.zeros(total, device=dev, dtype=torch.float32)
        self.loss_slot = self.garena[0:1]
        self.hdr = HDR          # the update skips the header
        self.gviews = []
        self.zero_slices = []   # (offset, count) of the gradients backward accumulates into: biases and the 1x1 head
        with torch.no_grad():
            for i, p in enumerate(params):
                n = p.numel()
                view = self.arena[offs[i]:offs[i] + n].view_as(p)
                view.copy_(p)
                p.data = view           # parameters now live in the arena (state_dict unchanged)
                self.gviews.append(self.garena[offs[i]:offs[i] + n].view_as(p))
                if p.dim() == 1 or p.shape[0] == 1:
                    self.zero_slices.append((offs[i], n))
        self.total = total
        # buckets in the order backward produces them: backend+head, ctx, conv4_x, the rest of the frontend
        # (1.7 M floats + the loss slot: all that is left to exchange when backward ends)
        o_hi, o_back, o_ctx = offs[2 * 7], offs[engine.IDX_BACKEND], offs[engine.IDX_CTX]
        self.buckets = GradBuckets(self.garena, {"backend": (o_back, o_ctx), "ctx": (o_ctx, total),
                                                 "frontend_hi": (o_hi, o_back), "frontend": (0, o_hi)},
                                   group=process_group)
        self.first = True
        model._wcache.invalidate()
        if self.world > 1:  # identical initial weights on every rank (train.py:104-114 / DDP broadcast)
            dist.broadcast(self.arena, src=0, group=self.pg)
            # NCCL's all-reduce kernels run next to backward.  They cannot share an SM with this library's one-CTA-per-SM
            # kernels, so CAN_B200_DP_RESERVE_SMS = k makes those kernels launch on SMs - k.  Default 0: measured at 2 GPUs
            # (three interleaved 40-step runs each) 8.53 ms per step with k = 0, 8.56 ms with k = 4 -- NCCL takes more than
            # a handful of SMs, and giving up enough of them costs more than the interference.
            if dev.type == "cuda" and "CAN_B200_CTAS" not in os.environ:
                reserve = int(os.environ.get("CAN_B200_DP_RESERVE_SMS", "0"))
                sms = torch.cuda.get_device_properties(dev).multi_processor_count
                if 0 < reserve < sms:
                    os.environ["CAN_B200_CTAS"] = str(sms - reserve)
        # Single-GPU steps of a fixed input shape are replayed from a CUDA graph after two eager steps: a step is
        # ~120 kernel launches (~2.4 ms of host time), which a loop that reads the loss back every step would
        # otherwise pay on the critical path.  CAN_B200_GRAPH=0 keeps every step eager.
        # Data parallel: NCCL calls are not captured (capturing them hung here); the step is captured as four graphs
        # (forward+loss+backend backward | contextual backward | frontend backward | SGD) sharing one memory pool,
        # and the bucketed all-reduces are issued eagerly between the replays.
        self.use_graph = os.environ.get("CAN_B200_GRAPH", "1") != "0" and (
            self.world == 1 or (dist.get_backend(self.pg) == "nccl" and os.environ.get("CAN_B200_GRAPH_DP", "1") != "0"))
        self._graphs = GraphCache()   # (img shape, gt shape) -> [graph(s), static img, static gt, static loss]
        self._captured_hyper = (self.lr, self.momentum)

    def step(self, img: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
        """One training step; returns the rank-mean of the local sum-losses as a device tensor [1] (a fresh tensor every
        call: it stays valid after later steps).  ``lr`` / ``momentum`` may be changed between steps (a captured step
        has them baked in as kernel arguments, so a change drops the captured graphs)."""
        if (self.lr, self.momentum) != self._captured_hyper:
            self._graphs = GraphCache()
            self._captured_hyper = (self.lr, self.momentum)
        if not self.use_graph or img.dtype != torch.float32 or gt.dtype != torch.float32:
            return self._step_eager(img, gt)
        key = (tuple(img.shape), tuple(gt.shape))
        ent = self._graphs.lookup(key)
        if ent is None:
            if not self._graphs.should_capture() or self.first:
                return self._step_eager(img, gt)
            try:
                xs, gs = img.clone(), gt.clone()
                torch.cuda.synchronize()
                if self.world == 1:
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        loss = self._step_eager(xs, gs)
                else:
                    graph, loss = self._capture_dp(xs, gs)
                ent = [graph, xs, gs, loss]
                self._graphs.insert(key, ent)
            except Exception as exc:  # noqa: BLE001 -- capture is an optimisation; eager stays correct
                print(f"can_b200: CUDA graph capture of the training step failed ({exc}); staying eager", file=sys.stderr)
                self.use_graph = False
                torch.cuda.synchronize()
                return self._step_eager(img, gt)
        graph, xs, gs, loss = ent
        xs.copy_(img, non_blocking=True)
        gs.copy_(gt, non_blocking=True)
        if self.world == 1:
            with nvtx.span("can.step.graph"):
                graph.replay()
            loss = loss.clone()   # the static tensor is rewritten by the next replay
        else:
            for g, tag in zip(graph[:-1], self.DP_TAGS):
                with nvtx.span("can.dp." + tag):
                    g.replay()
                    self.buckets.ready(tag)
            self.buckets.wait()
            loss = self.loss_slot / self.world   # rank-mean of the local sum-losses (came with the last bucket)
            with nvtx.span("can.sgd"):
                graph[-1].replay()
        ops.LAUNCHES += self._graph_launches
        self.model._wcache.invalidate()   # the replay updated the arena: packed weights are stale for eager callers
        return loss

    DP_TAGS = ("backend", "ctx", "frontend_hi", "frontend")   # the order engine.backward_steps yields them

    def _capture_dp(self, xs, gs):
        """One graph per gradient bucket + one for the SGD update (no NCCL inside); nothing executes during capture."""
        m = self.model
        launches0 = ops.LAUNCHES
        graphs = [torch.cuda.CUDAGraph() for _ in range(len(self.DP_TAGS) + 1)]
        pool = None
        gen = None
        for i, tag in enumerate(self.DP_TAGS):
            with torch.cuda.graph(graphs[i], pool=pool):
                if i == 0:
                    ops.zero_ranges(self.garena, self.zero_slices)
                    out, saved = engine.forward(m, xs, m.nsplit, save=True)
                    loss, _, _, grad = ops.mse_count(out, gs, want_grad=True, loss_out=self.loss_slot)
                    gen = engine.backward_steps(m, saved, grad, grads=self.gviews)
                assert next(gen) == tag
                if i == len(self.DP_TAGS) - 1:
                    for _ in gen:   # runs the generator to its end (nothing left to enqueue)
                        pass
            pool = graphs[0].pool()
        with torch.cuda.graph(graphs[-1], pool=pool):
            ops.sgd_momentum_step(self.arena[self.hdr:], self.garena[self.hdr:], self.buf[self.hdr:], self.lr,
                                  self.momentum, self.buckets.grad_scale, first=False)
        self._graph_launches = ops.LAUNCHES - launches0
        m._wcache.invalidate()
        return graphs, self.loss_slot

    def _step_eager(self, img: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
        m = self.model
        launches0 = ops.LAUNCHES
        # bias / head gradients are accumulated with atomics: those slices start at zero; every other gradient is
        # overwritten by its kernel (the arena was zeroed once, so the alignment padding stays zero)
        ops.zero_ranges(self.garena, self.zero_slices)
        out, saved = engine.forward(m, img, m.nsplit, save=True)
        # data parallel: the loss goes into the header slot of the gradient arena and is summed over ranks with the last
        # bucket (logged loss = rank-mean of the local sum-losses, reduce_value)
        with nvtx.span("can.loss"):
            loss, _, _, grad = ops.mse_count(out, gt, want_grad=True, loss_out=self.loss_slot if self.world > 1 else None)
        engine.backward(m, saved, grad, grads=self.gviews, on_ready=self.buckets.ready)
        if self.world > 1:
            self.buckets.wait()
            loss = self.loss_slot / self.world
        with nvtx.span("can.sgd"):
            ops.sgd_momentum_step(self.arena[self.hdr:], self.garena[self.hdr:], self.buf[self.hdr:], self.lr,
                                  self.momentum, self.buckets.grad_scale, first=self.first)
        self.first = False
        m._wcache.invalidate()  # the arena changed in place: packed weights must be rebuilt
        self._graph_launches = ops.LAUNCHES - launches0   # kernels of one step (what a graph replay launches)
        return loss


def train_one_epoch(model, optimizer, train_loader, device, epoch):
    """utils/train_eval_utils.py:13-58.  ``optimizer`` may be a FusedTrainer (fused path) or any
    torch optimizer (stock path through the autograd.Function).  Returns the mean loss."""
    model.train()
    mean_loss = torch.zeros(1, device=device)
    if is_main_process():
        try:
            from tqdm import tqdm
            train_loader = tqdm(train_loader, file=sys.stdout)
        except Exception:
            pass
    fused = isinstance(optimizer, FusedTrainer)
    # the next batch's host->device copies run on a side stream while this step computes (pinned loaders overlap fully)
    batches = DevicePrefetcher(train_loader, device) if torch.device(device).type == "cuda" else train_loader
    # Non-finite guard (train_eval_utils.py:48-50): the reference reads the loss back on every rank before
    # optimizer.step, a host sync per step.  Here step k's finite flag is copied to pinned host memory asynchronously and
    # looked at while step k+1 is being enqueued, so a diverged run stops one step late instead of one epoch late and no
    # step waits for the device.  (The late step's update is applied; the reference's exit leaves no usable state either.)
    on_cuda = torch.device(device).type == "cuda"
    flags = [None, None]   # (pinned bool [1], event) of the two steps in flight

    def check_late(slot):
        ent = flags[slot]
        if ent is not None:
            ent[1].synchronize()
            if not bool(ent[0][0]):
                print("WARNING: non-finite loss, ending training ", mean_loss)
                sys.exit(1)

    for step, (img, gt_dmap) in enumerate(batches):
        img = img.to(device, non_blocking=True)
        gt_dmap = gt_dmap.to(device, non_blocking=True)
        if on_cuda:
            check_late((step + 1) & 1)   # the flag of step - 1 (its kernels finished while this batch was fetched)
        nvtx.push("can.step")
        if fused:
            loss = optimizer.step(img, gt_dmap)
        else:
            optimizer.zero_grad()
            et_dmap = model(img)
            loss = torch.nn.functional.mse_loss(et_dmap, gt_dmap, reduction="sum")
            loss.backward()
            loss = reduce_value(loss.detach().reshape(1), average=True)
            optimizer.step()
        nvtx.pop()
        mean_loss = (mean_loss * step + loss.detach()) / (step + 1)
        if on_cuda:
            ent = flags[step & 1]
            if ent is None:
                ent = flags[step & 1] = (torch.empty(1, dtype=torch.bool).pin_memory(), torch.cuda.Event())
            ent[0].copy_(torch.isfinite(loss.detach()).reshape(-1)[:1], non_blocking=True)
            ent[1].record()
        if is_main_process() and hasattr(train_loader, "desc"):
            train_loader.desc = "[epoch {}]".format(epoch)   # (the running mean stays on the device: no per-step read-back)
    if not torch.isfinite(mean_loss).all():  # covers the last step (and CPU runs)
        print("WARNING: non-finite loss, ending training ", mean_loss)
        sys.exit(1)
    if torch.device(device).type == "cuda":   # (`device` may be a string, as in the reference's CLI)
        torch.cuda.synchronize(device)
    return mean_loss.item()


@torch.no_grad()
def evaluate(model, test_loader, device, epoch=0, show_images=False, use_wandb=False):
    """utils/train_eval_utils.py:61-138 without the plotting: returns sum |count(et) - count(gt)|
    over this rank's samples, reduced (summed) over ranks.  Counts come from the fused kernel, per
    image; like the reference (:83) the absolute difference is taken over the whole batch tensor."""
    model.eval()
    mae = torch.zeros(1, device=device)
    batches = DevicePrefetcher(test_loader, device) if torch.device(device).type == "cuda" else test_loader
    for img, gt_dmap in batches:
        img = img.to(device, non_blocking=True)
        gt_dmap = gt_dmap.to(device, non_blocking=True)
        et = model(img)
        _, ce, cg, _ = ops.mse_count(et, gt_dmap, want_grad=False)
        mae += (ce.sum() - cg.sum()).abs()
    if torch.device(device).type == "cuda":   # (`device` may be a string, as in the reference's CLI)
        torch.cuda.synchronize(device)
    mae = reduce_value(mae, average=False)
    return mae.item()


class DevicePrefetcher:
    """Pinned-host -> device copies of the NEXT batch on a side stream while the current step computes
    (the reference copies synchronously from pageable memory, utils/train_eval_utils.py:30-31).

    ``for img, gt in DevicePrefetcher(loader, device): ...`` yields device tensors whose copies have
    completed with respect to the current stream.  The copies land in two resident sets of device buffers
    used in turn (re-made only when a batch changes shape), so the steady state allocates nothing.  A
    yielded batch is valid until the NEXT batch is requested (its slot is refilled at the start of the iteration after
    that, ordered only after the work enqueued before that request): clone what must outlive the loop body."""

    _streams: dict = {}   # one copy stream per device for the life of the process

    def __init__(self, loader, device):
        self.loader, self.device = loader, torch.device(device)
        key = (self.device.type, self.device.index if self.device.index is not None else torch.cuda.current_device())
        if key not in DevicePrefetcher._streams:
            DevicePrefetcher._streams[key] = torch.cuda.Stream(device=self.device)
        self.stream = DevicePrefetcher._streams[key]
        self._bufs = [None, None]
        self._copied = [torch.cuda.Event(), torch.cuda.Event()]
        self._freed = [None, None]

    def _issue(self, batch, slot):
        bufs = self._bufs[slot]
        with torch.cuda.stream(self.stream):
            if bufs is None or len(bufs) != len(batch) or any(b.shape != t.shape or b.dtype != t.dtype
                                                              for b, t in zip(bufs, batch)):
                bufs = tuple(torch.empty(t.shape, dtype=t.dtype, device=self.device) for t in batch)
                self._bufs[slot] = bufs
            if self._freed[slot] is not None:
                self.stream.wait_event(self._freed[slot])   # the step that read this slot two batches ago
            for b, t in zip(bufs, batch):
                b.copy_(t, non_blocking=True)
            self._copied[slot].record(self.stream)
        return bufs

    def __iter__(self):
        it = iter(self.loader)
        try:
            pending = self._issue(next(it), 0)
        except StopIteration:
            return
        slot = 0
        while pending is not None:
            dev = pending
            try:
                pending = self._issue(next(it), slot ^ 1)   # next batch's H2D overlaps this batch's step
            except StopIteration:
                pending = None
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(self._copied[slot])
            for t in dev:
                t.record_stream(cur)
            try:
                yield dev
            finally:
                # the consumer has enqueued its step (or left the loop early): the slot may be overwritten once the
                # work enqueued so far is done
                if self._freed[slot] is None:
                    self._freed[slot] = torch.cuda.Event()
                self._freed[slot].record(torch.cuda.current_stream(self.device))
            slot ^= 1
