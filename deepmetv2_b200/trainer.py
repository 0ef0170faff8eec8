"""Fused, sync-free training / inference step for GraphMETNetwork (the body of train.py:39-55 and
evaluate.py:81-100 of the reference as ONE call).

``GraphMETTrainer.train_step(batch)`` enqueues, on the current stream and without any host sync:
    etaphi prep -> radius graph -> fused forward (train-mode BN) -> per-event MET + loss (+ dL/dw)
    -> fused backward into a flat gradient buffer -> [NCCL all-reduce of the 26.6 KB buffer] -> fused flat AdamW.
The radius graph is, by default, kept IMPLICIT (``graph="implicit"``, ``gm_geo_build``): points are cell-binned,
the network runs in cell-sorted node order and the EdgeConv kernels re-evaluate the edge predicate on cell
neighbourhoods staged in shared memory, so no edge list and no transposed graph are ever written.
``graph="csr"`` is the explicit path (count, scan, fill; E stays on the device; transposed CSR iff capped).  The
tcgen05 ``edge_precision`` variants (tf32 / bf16) run the per-edge contraction on the implicit graph too
(``csrc/geo_tc.cu``), in eval mode and -- for ``aggr="max"``, with the arg-max sources recorded for the backward --
in training.

Data parallelism (``process_group=``): every rank steps on its own events; the per-event-SUM gradients are all-reduced
and divided by the GLOBAL event count, so ranks may hold different numbers of events (a rank with none still joins
with a zero gradient) and the update equals the gradient of the mean loss over all events of the step.  The exchange
is fused with AdamW in one kernel over NVLink peer memory (``gm_allreduce_adamw_step``, ``csrc/comm.cu``); if CUDA IPC
is unavailable it falls back to ``torch.distributed.all_reduce`` + ``gm_adamw_step_ex``.  BatchNorm uses rank-local
batch statistics (DistributedDataParallel's default); the running statistics are rank-local too and
:meth:`sync_bn_buffers` makes rank 0's authoritative before a checkpoint / evaluation (what DDP's buffer broadcast does).
Parameters stay the module's ``nn.Parameter``s (re-pointed at one flat fp32 buffer), so
``state_dict()`` / checkpoints are unchanged; optimizer state converts to and from the
``torch.optim.AdamW`` layout the reference's checkpoints store (``optim_dict``).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib
from ._lib import AGGR, PRECISION, GmAdamW, check, lib, ptr, stream
from .synth import EventBatch

_FUSED_ADAMW = os.environ.get("GM_FUSED_ADAMW", "1") != "0"     # 0: separate gm_net_backward_geo + gm_adamw_step_ex launches


class FusedAdamW(torch.optim.Optimizer):
    """``torch.optim.AdamW``-shaped handle on the trainer's fused optimiser (train.py:75-76): hyper-parameters live in
    ``param_groups[0]`` exactly as in torch, so ``torch.optim.lr_scheduler.ReduceLROnPlateau(trainer.optimizer, factor=0.5,
    patience=500, threshold=0.05)`` / ``scheduler.step(loss)`` (train.py:58,76) work unchanged -- the fused step kernel
    reads ``lr`` / ``betas`` / ``eps`` / ``weight_decay`` from the group on every step.  ``step()`` is a no-op (the update
    is the last kernel of ``GraphMETTrainer.train_step``); ``state_dict()`` / ``load_state_dict()`` use torch's AdamW
    layout (what the reference stores as ``optim_dict``)."""

    def __init__(self, trainer, params, lr, betas, eps, weight_decay):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False))
        self._trainer = trainer

    def step(self, closure=None):          # the fused kernel already applied the update
        return None if closure is None else closure()

    def zero_grad(self, set_to_none: bool = True):   # gradients are overwritten by every backward
        return None

    def state_dict(self):
        return self._trainer.optimizer_state_dict()

    def load_state_dict(self, sd):
        self._trainer.load_optimizer_state_dict(sd)


class GraphMETTrainer:
    def __init__(self, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01, r=0.4,
                 max_num_neighbors=255, loop=True, phi_wrap=False, process_group=None, device=None,
                 graph="implicit", fused_allreduce=True, knn_k=20, overlap_exchange=True):
        self.model = model
        self.net = model.graphnet if hasattr(model, "graphnet") else model
        self.apply_sigmoid = hasattr(model, "graphnet")       # Net applies the sigmoid, GraphMETNetwork does not
        self.r, self.max_nn, self.loop, self.phi_wrap = r, max_num_neighbors, loop, phi_wrap
        if graph not in ("implicit", "csr", "knn"):
            raise ValueError("graph must be 'implicit', 'csr' or 'knn'")
        self.graph = graph        # "knn": the per-layer dynamic graph of model/graph_met_network.py:63 (row f4)
        self.knn_k = int(knn_k)
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        self.rank = torch.distributed.get_rank(process_group) if process_group is not None else 0
        self.step_count = 0
        params = list(self.net.parameters())
        dev = device or params[0].device
        if dev.type != "cuda":
            raise _lib.GraphMETLibraryError("GraphMETTrainer needs the model on a CUDA device (no CPU fallback)")
        self.device = dev
        n = sum(p.numel() for p in params)
        self.flat = torch.empty(n, dtype=torch.float32, device=dev)
        # [gradient | local events | local events * local loss]: the tail rides along in the all-reduce fallback
        self._payload = torch.zeros(n + 2, dtype=torch.float32, device=dev)
        self.flat_grad = self._payload[:n]
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self._views, self._gviews, off = [], [], 0
        for p in params:
            k = p.numel()
            self.flat[off:off + k].copy_(p.data.reshape(-1))
            p.data = self.flat[off:off + k].view_as(p)        # parameters now alias the flat buffer
            self._views.append(p.data)
            self._gviews.append(self.flat_grad[off:off + k].view_as(p))
            off += k
        self._P = self.net._param_struct()
        self._G = self.net._param_struct(grad_views=self._gviews)
        self.optimizer = FusedAdamW(self, params, lr, tuple(betas), eps, weight_decay)
        self._cap_nodes = self._cap_events = self._cap_edges = -1
        self._cap_implicit = None
        self.flags = torch.zeros(1, dtype=torch.int32, device=dev)
        self.stats = torch.zeros(4, dtype=torch.int32, device=dev)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)          # local mean loss of the last step
        self.global_loss = torch.zeros((), dtype=torch.float32, device=dev)   # mean over ALL ranks' events (DP)
        self._host_loss = torch.zeros((), dtype=torch.float32).pin_memory()
        # {step, forward error flags, graph-build flags} mirrored by the step's last kernel into pinned host memory
        self._status = torch.zeros(4, dtype=torch.int32).pin_memory()
        self._status_seen = 0
        self.comm = None
        # overlap_exchange (peer-memory exchange only): a step ENDS by pushing its gradient to the peers and the wait +
        # AdamW run at the start of the next step, behind that step's radius-graph build (which does not read the
        # weights), so rank skew up to one graph build is hidden.  flush() applies a pending update at once.
        self.overlap_exchange = bool(overlap_exchange)
        self._pending = None
        if self.world > 1 and fused_allreduce:
            self._connect_peers(n)

    # hyper-parameters: one source of truth, the optimizer's param group (so LR schedulers act on the fused step)
    def _hp(self, key):
        return self.optimizer.param_groups[0][key]

    lr = property(lambda self: self._hp("lr"), lambda self, v: self.optimizer.param_groups[0].__setitem__("lr", v))
    betas = property(lambda self: tuple(self._hp("betas")),
                     lambda self, v: self.optimizer.param_groups[0].__setitem__("betas", tuple(v)))
    eps = property(lambda self: self._hp("eps"), lambda self, v: self.optimizer.param_groups[0].__setitem__("eps", v))
    weight_decay = property(lambda self: self._hp("weight_decay"),
                            lambda self, v: self.optimizer.param_groups[0].__setitem__("weight_decay", v))

    # ------------------------------------------------------------------------------------------ data parallelism
    def _connect_peers(self, n: int):
        """Exchange CUDA IPC handles of the per-rank exchange buffers (csrc/comm.cu).  Every rank must agree: if any
        rank cannot create or map a buffer, all fall back to torch.distributed.all_reduce."""
        import torch.distributed as dist
        l = lib()
        comm = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        with torch.cuda.device(self.device):
            ok = l.gm_comm_create(n, self.rank, self.world, C.byref(comm), handle) == _lib.GM_OK
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle) if ok else None, group=self.pg)
            if ok and all(h is not None for h in handles):
                ok = l.gm_comm_connect(comm, b"".join(handles)) == _lib.GM_OK
            else:
                ok = False
            oks = [None] * self.world
            dist.all_gather_object(oks, bool(ok), group=self.pg)     # also: nobody steps before everybody has mapped
        if all(oks):
            self.comm = comm
        elif comm:
            l.gm_comm_destroy(comm)

    def flush(self):
        """Apply the gradient exchange of the last step if it is still pending (``overlap_exchange``): parameters,
        optimizer state and ``global_loss`` are up to date afterwards.  Every rank must call it at the same point."""
        if self._pending is not None:
            hyper, self._pending = self._pending, None
            check(lib().gm_exchange_apply_adamw(ptr(self.flat), ptr(self.exp_avg), ptr(self.exp_avg_sq), self.flat.numel(),
                                                *hyper, self.comm, ptr(self.global_loss), ptr(self.flags), ptr(self.stats),
                                                ptr(self._status), stream(self.device)), "gm_exchange_apply_adamw")

    def close(self):
        if self.comm is not None:
            self.flush()
            lib().gm_comm_destroy(self.comm)
            self.comm = None

    def __del__(self):
        try:
            self.close()
        except Exception:      # interpreter shutdown
            pass

    def sync_bn_buffers(self, src: int = 0):
        """BatchNorm running statistics are rank-local during DP training; make rank ``src``'s authoritative (call before
        saving a checkpoint or evaluating -- the buffer broadcast DistributedDataParallel does every forward)."""
        if self.world == 1:
            return
        self.flush()
        import torch.distributed as dist
        for _, buf in self.net.named_buffers():
            dist.broadcast(buf, src=dist.get_global_rank(self.pg, src), group=self.pg)

    # ------------------------------------------------------------------------------------------ deferred errors
    def _raise_pending(self):
        """Device-side error flags of an EARLIER step (mirrored into pinned memory by that step's last kernel): raise
        one step late, like a CUDA device assert, without a copy or a sync on the hot path."""
        st = self._status
        step = int(st[0])
        if step == self._status_seen:
            return
        self._status_seen = step
        fwd, graph = int(st[1]), int(st[2])
        if fwd & _lib.FLAG_CAT_OUT_OF_RANGE:
            self.flags.zero_()
            raise IndexError("index out of range in self (embed_charge / embed_pdgid / embed_pv)")
        if graph & _lib.FLAG_EVENT_TOO_LARGE:
            raise _lib.GraphMETLibraryError("an event has more nodes than max_event_nodes promised: the radius graph "
                                            f"of step {step} is wrong")

    def check_errors(self):
        """Synchronising check of the device error flags (use after eval_step, or at the end of an epoch)."""
        flags = int(self.flags.item())
        graph = int(self.stats[3].item()) if self._cap_nodes >= 0 else 0
        if flags & _lib.FLAG_CAT_OUT_OF_RANGE:
            self.flags.zero_()
            raise IndexError("index out of range in self (embed_charge / embed_pdgid / embed_pv)")
        if graph & _lib.FLAG_EVENT_TOO_LARGE:
            raise _lib.GraphMETLibraryError("an event has more nodes than max_event_nodes promised")

    def _check_batch(self, x, y, p):
        if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] < 11 or x.stride(1) != 1:
            raise ValueError("batch.x must be float32 [N, >=11] with unit column stride")
        if y.dtype != torch.float32 or y.dim() != 2 or y.shape[1] < 2 or (y.shape[0] and y.stride(1) != 1):
            raise ValueError("batch.y must be float32 [B, >=2] with unit column stride")
        if p.dtype != torch.int32 or p.dim() != 1 or p.numel() != y.shape[0] + 1 or not p.is_contiguous():
            raise ValueError("batch.ptr must be a contiguous int32 [B+1] tensor")
        if x.device != self.device or y.device != self.device or p.device != self.device:
            raise ValueError(f"batch tensors must live on {self.device}")

    # ------------------------------------------------------------------------------------------ buffers
    def _implicit(self, training: bool | None = None) -> bool:
        """fp32: always; tf32 / bf16 (per-edge tcgen05 on the implicit graph, csrc/geo_tc.cu): eval, and training with
        max aggregation (the kernel records the arg-max sources the backward routes gradients through)."""
        if training is None:
            training = self.net.training
        if self.graph != "implicit":
            return False
        return self.net.edge_precision == "fp32" or \
            (self.net.edge_precision in ("tf32", "bf16") and (not training or self.net.aggr == "max"))

    def _reserve(self, N: int, B: int, max_event_nodes: int):
        implicit = self._implicit()
        knn = self.graph == "knn"
        E = 0 if (implicit or knn) else N * min(self.max_nn + (0 if self.loop else 1), max(max_event_nodes, 1))
        if N <= self._cap_nodes and B <= self._cap_events and E <= self._cap_edges and \
                implicit == self._cap_implicit:
            return
        N, B, E = max(N, self._cap_nodes), max(B, self._cap_events), max(E, self._cap_edges)
        l, dev, depth = lib(), self.device, len(self.net.conv_continuous)
        u8 = lambda nbytes: torch.empty(int(nbytes), dtype=torch.uint8, device=dev)  # noqa: E731
        self.etaphi = torch.empty(N, 2, dtype=torch.float32, device=dev)
        if implicit:      # no edge list: O(N) state only
            self.geo_ws = u8(l.gm_geo_ws_bytes(N, B))
            self.perm = torch.empty(max(N, 1), dtype=torch.int32, device=dev)
        elif knn:         # per-layer k-NN CSR + transposes
            self.knn_ws = u8(l.gm_net_knn_ws_bytes(N, depth, self.knn_k))
        else:
            self.rg_ws = u8(l.gm_radius_graph_ws_bytes(N, B))
            self.rowptr = torch.empty(N + 1, dtype=torch.int32, device=dev)
            self.col = torch.empty(max(E, 1), dtype=torch.int32, device=dev)
            self.rowptr_t = torch.empty(N + 1, dtype=torch.int32, device=dev)
            self.col_t = torch.empty(max(E, 1), dtype=torch.int32, device=dev)
            self.csr_ws = u8(l.gm_csr_ws_bytes(N, E))
        self.fwd_ws = u8(l.gm_net_ws_bytes(N, depth, 1))
        self.bwd_ws = u8(l.gm_net_bwd_ws_bytes(N, depth))
        self.bwd_ws[:256].zero_()           # "last CTA" tickets: gm_net_backward_geo_adamw wants them zero on first use
        self.weights = torch.empty(N, dtype=torch.float32, device=dev)
        self.grad_w = torch.empty(N, dtype=torch.float32, device=dev)
        self.met = torch.empty(B, 2, dtype=torch.float32, device=dev)
        self._cap_nodes, self._cap_events, self._cap_edges, self._cap_implicit = N, B, E, implicit

    # ------------------------------------------------------------------------------------------ pieces
    def _graph(self, x, ptr_i32, N, B, max_event_nodes, need_transpose):
        if self.graph == "knn":      # the graphs are built layer by layer inside the forward
            self._knn_ptr, self._knn_B = ptr_i32, B
            return
        l, st = lib(), stream(self.device)
        check(l.gm_prepare_etaphi(ptr(x), x.stride(0), N, ptr(self.etaphi), st), "gm_prepare_etaphi")
        if self._implicit():
            check(l.gm_geo_build(ptr(self.etaphi), 2, 2, ptr(ptr_i32), N, B, float(self.r), int(self.max_nn),
                                 int(self.loop), int(self.phi_wrap), int(max_event_nodes), ptr(self.geo_ws),
                                 self.geo_ws.numel(), ptr(self.perm), ptr(self.stats), st), "gm_geo_build")
            return
        args = (ptr(self.etaphi), 2, 2, ptr(ptr_i32), N, B, float(self.r), int(self.max_nn), int(self.loop),
                int(self.phi_wrap), int(max_event_nodes), ptr(self.rg_ws), self.rg_ws.numel())
        check(l.gm_radius_graph_count(*args, ptr(self.rowptr), ptr(self.stats), st), "gm_radius_graph_count")
        check(l.gm_radius_graph_fill(*args, ptr(self.rowptr), ptr(self.col), self._cap_edges, None,
                                     ptr(self.stats), st), "gm_radius_graph_fill")
        if need_transpose:   # kernels are no-ops unless some node hit the cap (decided on the device)
            check(l.gm_csr_transpose(ptr(self.rowptr), ptr(self.col), N, self._cap_edges, ptr(self.csr_ws),
                                     self.csr_ws.numel(), ptr(self.rowptr_t), ptr(self.col_t),
                                     self.stats.data_ptr() + 8, st), "gm_csr_transpose")

    def _forward(self, x, N, B, training, save):
        net = self.net
        if self.graph == "knn":
            if net.edge_precision != "fp32":
                raise NotImplementedError("graph='knn' runs the fp32 EdgeConv")
            check(lib().gm_net_forward_knn(C.byref(self._P), ptr(x), x.stride(0), None, x.data_ptr() + 8 * 4, x.stride(0),
                                           ptr(self._knn_ptr), N, B, self.knn_k, int(training), int(save),
                                           int(self.apply_sigmoid), AGGR[net.aggr], float(net.bn_all.momentum),
                                           float(net.bn_all.eps), ptr(self.fwd_ws), self.fwd_ws.numel(), ptr(self.knn_ws),
                                           self.knn_ws.numel(), ptr(self.weights), ptr(self.flags), stream(self.device)),
                  "gm_net_forward_knn")
            return
        if self._implicit():
            check(lib().gm_net_forward_geo(C.byref(self._P), ptr(x), x.stride(0), None, x.data_ptr() + 8 * 4,
                                           x.stride(0), ptr(self.geo_ws), N, B, ptr(self.perm), int(training),
                                           int(save), int(self.apply_sigmoid), AGGR[net.aggr],
                                           PRECISION[net.edge_precision], float(net.bn_all.momentum), float(net.bn_all.eps), ptr(self.fwd_ws),
                                           self.fwd_ws.numel(), ptr(self.weights), ptr(self.flags), stream(self.device)),
                  "gm_net_forward_geo")
            return
        check(lib().gm_net_forward(C.byref(self._P), ptr(x), x.stride(0), None, x.data_ptr() + 8 * 4, x.stride(0),
                                   ptr(self.rowptr), ptr(self.col), N, None, int(training), int(save),
                                   int(self.apply_sigmoid), AGGR[net.aggr], PRECISION[net.edge_precision],
                                   float(net.bn_all.momentum), float(net.bn_all.eps), ptr(self.fwd_ws),
                                   self.fwd_ws.numel(), ptr(self.weights), ptr(self.flags), stream(self.device)),
              "gm_net_forward")

    def _backward(self, x, N, B):
        net, l, st = self.net, lib(), stream(self.device)
        if self.graph == "knn":
            check(l.gm_net_backward_knn(C.byref(self._P), C.byref(self._G), ptr(x), x.stride(0), None, x.data_ptr() + 32,
                                        x.stride(0), N, self.knn_k, 1, int(self.apply_sigmoid), AGGR[net.aggr],
                                        ptr(self.fwd_ws), self.fwd_ws.numel(), ptr(self.bwd_ws), self.bwd_ws.numel(),
                                        ptr(self.knn_ws), self.knn_ws.numel(), ptr(self.weights), ptr(self.grad_w), st),
                  "gm_net_backward_knn")
            return
        if self._implicit():
            check(l.gm_net_backward_geo(C.byref(self._P), C.byref(self._G), ptr(x), x.stride(0), None,
                                        x.data_ptr() + 32, x.stride(0), ptr(self.geo_ws), N, B, ptr(self.perm), 1,
                                        int(self.apply_sigmoid), AGGR[net.aggr], ptr(self.fwd_ws), self.fwd_ws.numel(),
                                        ptr(self.bwd_ws), self.bwd_ws.numel(), ptr(self.weights), ptr(self.grad_w), st),
                  "gm_net_backward_geo")
            return
        check(l.gm_net_backward(C.byref(self._P), C.byref(self._G), ptr(x), x.stride(0), None, x.data_ptr() + 32,
                                x.stride(0), ptr(self.rowptr), ptr(self.col), ptr(self.rowptr_t), ptr(self.col_t),
                                self.stats.data_ptr() + 8, N, None, 1, int(self.apply_sigmoid), AGGR[net.aggr],
                                float(net.bn_all.eps), ptr(self.fwd_ws), self.fwd_ws.numel(), ptr(self.bwd_ws),
                                self.bwd_ws.numel(), ptr(self.weights), ptr(self.grad_w), st), "gm_net_backward")

    @staticmethod
    def _sizes(batch: EventBatch, max_event_nodes):
        N, B = batch.x.shape[0], batch.y.shape[0]
        if max_event_nodes is None:
            max_event_nodes = getattr(batch, "max_event_nodes", None)
        if max_event_nodes is None:
            max_event_nodes = int((batch.ptr[1:] - batch.ptr[:-1]).max().item()) if B else 0   # syncs; pass it in
        return N, B, int(max_event_nodes)

    # ------------------------------------------------------------------------------------------ public
    def train_step(self, batch: EventBatch, max_event_nodes: int | None = None) -> torch.Tensor:
        """One optimisation step on a device-resident batch. Returns the loss as a 0-dim device tensor
        (no sync): the mean over this rank's events, or, with a process group, over ALL ranks' events.
        ``batch.x`` [N,11] f32, ``batch.y`` [B,>=2] f32, ``batch.ptr`` [B+1] i32.  B == 0 is allowed (a rank whose
        shard is empty still takes part in the gradient exchange, contributing zero)."""
        self._raise_pending()
        x, y, p = batch.x, batch.y, batch.ptr
        self._check_batch(x, y, p)
        N, B, men = self._sizes(batch, max_event_nodes)
        if not self.net.training:      # nn.Module.train() walks the module tree: not once per step
            self.net.train()
        l, st = lib(), stream(self.device)
        dp = self.world > 1
        self._last_N = N
        if N > 0 and B > 0:
            self._reserve(N, B, men)
            self._graph(x, p, N, B, men, need_transpose=True)
            self.flush()            # the previous step's update, behind this step's (weight-independent) graph build
            self._forward(x, N, B, training=True, save=True)
            # DP: gradient of the per-event SUM; the optimiser kernel divides by the global event count
            check(l.gm_met_loss_ex(ptr(self.weights), ptr(x), x.stride(0), ptr(y), y.stride(0), ptr(p), N, B,
                                   1.0 if dp else 1.0 / B, ptr(self.met), ptr(self.loss), ptr(self.grad_w), st),
                  "gm_met_loss")
            if not dp and self.graph != "knn" and self._implicit() and _FUSED_ADAMW:
                # single GPU: loss.backward() + optimizer.step() as ONE call, AdamW in the last kernel of the backward
                self.step_count += 1
                opt = GmAdamW(ptr(self.flat), ptr(self.flat_grad), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                              self.flat.numel(), float(self.lr), float(self.betas[0]), float(self.betas[1]),
                              float(self.eps), float(self.weight_decay), self.step_count, 1.0, ptr(self.flags),
                              ptr(self.stats), ptr(self._status))
                check(l.gm_net_backward_geo_adamw(
                    C.byref(self._P), C.byref(self._G), ptr(x), x.stride(0), None, x.data_ptr() + 32, x.stride(0),
                    ptr(self.geo_ws), N, B, ptr(self.perm), 1, int(self.apply_sigmoid), AGGR[self.net.aggr],
                    ptr(self.fwd_ws), self.fwd_ws.numel(), ptr(self.bwd_ws), self.bwd_ws.numel(), ptr(self.weights),
                    ptr(self.grad_w), C.byref(opt), st), "gm_net_backward_geo_adamw")
                return self.loss
            self._backward(x, N, B)
        else:                       # empty shard: nothing to differentiate, but the exchange below must still happen
            self.flush()
            self._payload.zero_()
            self.loss.zero_()
            B = 0
        self.step_count += 1
        hyper = (float(self.lr), float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay),
                 self.step_count)
        mirrors = (ptr(self.flags), ptr(self.stats), ptr(self._status))
        if not dp:
            check(l.gm_adamw_step_ex(ptr(self.flat), ptr(self.flat_grad), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                                     self.flat.numel(), *hyper, 1.0, None, *mirrors, st), "gm_adamw_step")
            return self.loss
        if self.comm is not None and self.overlap_exchange:
            # push now, wait + AdamW at the start of the next step (or in flush()); global_loss lags by that much
            check(l.gm_exchange_post(ptr(self.flat_grad), self.flat.numel(), float(B), ptr(self.loss), self.comm, st),
                  "gm_exchange_post")
            self._pending = hyper
            return self.global_loss
        if self.comm is not None:   # ONE kernel: push to peers over NVLink, wait, rank-ordered sum, AdamW
            check(l.gm_allreduce_adamw_step(ptr(self.flat), ptr(self.flat_grad), ptr(self.exp_avg),
                                            ptr(self.exp_avg_sq), self.flat.numel(), *hyper, float(B), ptr(self.loss),
                                            self.comm, ptr(self.global_loss), *mirrors, st), "gm_allreduce_adamw_step")
            return self.global_loss
        n = self.flat.numel()       # fallback: NCCL SUM all-reduce of [gradient | events | events * loss]
        self._payload[n].fill_(float(B))
        torch.mul(self.loss, float(B), out=self._payload[n + 1])
        torch.distributed.all_reduce(self._payload, group=self.pg)
        check(l.gm_adamw_step_ex(ptr(self.flat), ptr(self.flat_grad), ptr(self.exp_avg), ptr(self.exp_avg_sq), n,
                                 *hyper, 1.0, self._payload.data_ptr() + 4 * n, *mirrors, st), "gm_adamw_step")
        torch.div(self._payload[n + 1], self._payload[n].clamp_min(1.0), out=self.global_loss)
        return self.global_loss

    def train_step_from_host(self, host_batch: EventBatch, max_event_nodes: int | None = None) -> float:
        """End-to-end step from PINNED host buffers: H2D copies, the step, and the loss read back
        (what train.py:41,54 do per batch)."""
        dev = self.device
        x = host_batch.x.to(dev, non_blocking=True)
        y = host_batch.y.to(dev, non_blocking=True)
        p = host_batch.ptr.to(dev, non_blocking=True)
        men = max_event_nodes if max_event_nodes is not None else \
            int((host_batch.ptr[1:] - host_batch.ptr[:-1]).max())       # host tensor: no device sync
        loss = self.train_step(EventBatch(x, y, None, p), men)
        self.flush()                # the caller reads this step's (global) loss: no deferral on this path
        self._host_loss.copy_(loss, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return float(self._host_loss)

    def _epoch_from_host(self, host_batches, depth, step):
        """Double-buffered H2D (``feeder.HostFeeder``) + loss read-back deferred by ONE step: the 4-byte D2H of step
        k is enqueued right behind its kernels, but the host waits for it only after step k+1 has been enqueued, so
        neither the copy engine nor the SMs ever wait for Python."""
        from .feeder import HostFeeder
        if not hasattr(self, "_loss_ring"):
            self._loss_ring = [(torch.zeros((), dtype=torch.float32).pin_memory(), torch.cuda.Event()) for _ in range(2)]
        losses, pending = [], None
        for k, b in enumerate(HostFeeder(host_batches, self.device, depth=depth)):
            loss = step(b)
            self.flush()            # every step's loss is read back: apply the exchange before the copy is enqueued
            slot, ev = self._loss_ring[k & 1]
            slot.copy_(loss, non_blocking=True)
            ev.record()
            if pending is not None:
                pending[1].synchronize()
                losses.append(float(pending[0]))
            pending = (slot, ev)
        if pending is not None:
            pending[1].synchronize()
            losses.append(float(pending[0]))
        return losses

    def train_epoch_from_host(self, host_batches, depth: int = 2):
        """train.py:39-55 over an iterable of PINNED host batches with the H2D copy of batch k+1 overlapped with
        the step on batch k; the loss of every step is read back like train.py:54 does (one step late).
        Returns the list of per-step losses."""
        return self._epoch_from_host(host_batches, depth, lambda b: self.train_step(b, b.max_event_nodes))

    def eval_epoch_from_host(self, host_batches, depth: int = 2):
        """evaluate.py:71-103 over pinned host batches, fed like :meth:`train_epoch_from_host`.  Returns the
        per-batch losses (weights / MET of the last batch stay in ``self.weights`` / ``self.met``)."""
        return self._epoch_from_host(host_batches, depth, lambda b: self.eval_step(b, b.max_event_nodes)[1])

    @torch.no_grad()
    def eval_step(self, batch: EventBatch, max_event_nodes: int | None = None):
        """evaluate.py:81-100 fused: returns (weights[N] view, loss 0-dim, met[B,2]) on the device."""
        self.flush()
        x, y, p = batch.x, batch.y, batch.ptr
        self._check_batch(x, y, p)
        N, B, men = self._sizes(batch, max_event_nodes)
        if self.net.training:
            self.net.eval()
        self._reserve(N, B, men)
        self._graph(x, p, N, B, men, need_transpose=False)
        self._forward(x, N, B, training=False, save=False)
        check(lib().gm_met_loss(ptr(self.weights), ptr(x), x.stride(0), ptr(y), y.stride(0), ptr(p), N, B,
                                ptr(self.met), ptr(self.loss), None, stream(self.device)), "gm_met_loss")
        return self.weights[:N], self.loss, self.met[:B]

    def knn_layer_graph(self, layer: int):
        """(rowptr [N+1], col [E]) int32 of the k-NN graph EdgeConv layer ``layer`` used in the last step (graph="knn";
        synchronises; tests / inspection).  Mirrors knn_layers_layout in csrc/net_layout.cuh."""
        N, k = self._last_N, self.knn_k
        al = lambda v: (v + 255) // 256 * 256  # noqa: E731
        per_layer = 2 * (al((N + 1) * 4) + al(N * k * 4))
        off = layer * per_layer
        rowptr = self.knn_ws[off:off + (N + 1) * 4].view(torch.int32).clone()
        off += al((N + 1) * 4)
        E = int(rowptr[-1])
        return rowptr, self.knn_ws[off:off + E * 4].view(torch.int32).clone()

    def graph_stats(self):
        """(E, max degree before the cap, #capped nodes, flags) of the last batch — synchronises.  On the
        implicit path E is -1 (edges are never counted) and the max degree covers cap candidates only."""
        return tuple(self.stats.tolist())

    # ------------------------------------------------------------------------------------------ optimizer state
    def optimizer_state_dict(self):
        """State in torch.optim.AdamW's layout (what the reference stores as ``optim_dict``)."""
        if getattr(self, "_pending", None) is not None:
            self.flush()
        state, off = {}, 0
        for i, v in enumerate(self._views):
            k = v.numel()
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self.exp_avg[off:off + k].view_as(v).clone(),
                        "exp_avg_sq": self.exp_avg_sq[off:off + k].view_as(v).clone()}
            off += k
        group = {"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay,
                 "amsgrad": False, "params": list(range(len(self._views)))}
        return {"state": state, "param_groups": [group]}

    def load_optimizer_state_dict(self, sd):
        g = sd["param_groups"][0]
        self.lr, self.betas, self.eps = g["lr"], tuple(g["betas"]), g["eps"]
        self.weight_decay = g["weight_decay"]
        off = 0
        for i, v in enumerate(self._views):
            k = v.numel()
            st = sd["state"].get(i)
            if st is not None:
                self.exp_avg[off:off + k].copy_(st["exp_avg"].reshape(-1))
                self.exp_avg_sq[off:off + k].copy_(st["exp_avg_sq"].reshape(-1))
                self.step_count = int(st["step"])
            off += k


class DRNTrainer:
    """Fused, sync-free training step for the DynamicReductionNetwork variant (BASELINE config 4, row f1):
    forward (``gm_drn_forward``, train-mode BatchNorm over the edge rows) -> wrapper activations + Cartesian MET loss
    (``gm_drn_met_loss``) -> backward (``gm_drn_backward``, all 31 gradients into one flat buffer) -> fused flat AdamW
    (``gm_adamw_step``).  Parameters stay the module's ``nn.Parameter``s (re-pointed at a flat fp32 buffer);
    ``trainer.optimizer`` is the same torch-Optimizer-shaped handle as GraphMETTrainer's.
    ``net.edge_precision = "bf16"`` runs the FORWARD's 96 -> 64 edge contraction on tcgen05 (``k_drn_edge_tc``); the
    backward stays fp32 and routes through the arg-max sources that forward recorded (the bf16 rounding is treated as
    the identity: a straight-through estimate, reported under a looser tolerance)."""

    def __init__(self, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01):
        self.model = model
        self.net = model.drn if hasattr(model, "drn") else model
        params = list(self.net.parameters())
        dev = params[0].device
        if dev.type != "cuda":
            raise _lib.GraphMETLibraryError("DRNTrainer needs the model on a CUDA device (no CPU fallback)")
        if self.net.output_dim != 2:
            raise NotImplementedError("the fused loss is the (MET, MET phi) regression of model/net.py:12-37: output_dim = 2")
        self.device = dev
        n = sum(p.numel() for p in params)
        self.flat = torch.empty(n, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self._views, gviews, off = [], {}, 0
        for (name, p) in self.net.named_parameters():
            k = p.numel()
            self.flat[off:off + k].copy_(p.data.reshape(-1))
            p.data = self.flat[off:off + k].view_as(p)
            self._views.append(p.data)
            gviews[name] = self.flat_grad[off:off + k].view_as(p)
            off += k
        self._gviews = gviews
        self._P = self.net._param_struct()
        self._G = self.net._param_struct(grad_views=gviews)
        self.optimizer = FusedAdamW(self, params, lr, tuple(betas), eps, weight_decay)
        self.step_count = 0
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self._host_loss = torch.zeros((), dtype=torch.float32).pin_memory()
        self._cap = (-1, -1)

    lr = property(lambda self: self.optimizer.param_groups[0]["lr"])

    def optimizer_state_dict(self):
        return GraphMETTrainer.optimizer_state_dict(self)

    def load_optimizer_state_dict(self, sd):
        g = sd["param_groups"][0]
        self.optimizer.param_groups[0].update(lr=g["lr"], betas=tuple(g["betas"]), eps=g["eps"], weight_decay=g["weight_decay"])
        off = 0
        for i, v in enumerate(self._views):
            k, st = v.numel(), sd["state"].get(i)
            if st is not None:
                self.exp_avg[off:off + k].copy_(st["exp_avg"].reshape(-1))
                self.exp_avg_sq[off:off + k].copy_(st["exp_avg_sq"].reshape(-1))
                self.step_count = int(st["step"])
            off += k

    betas = property(lambda self: tuple(self.optimizer.param_groups[0]["betas"]))
    eps = property(lambda self: self.optimizer.param_groups[0]["eps"])
    weight_decay = property(lambda self: self.optimizer.param_groups[0]["weight_decay"])

    def _reserve(self, N, B):
        if N <= self._cap[0] and B <= self._cap[1]:
            return
        N, B = max(N, self._cap[0]), max(B, self._cap[1])
        l, dev, k = lib(), self.device, int(self.net.k)
        self.fwd_ws = torch.empty(l.gm_drn_ws_bytes(N, B, k), dtype=torch.uint8, device=dev)
        self.bwd_ws = torch.empty(l.gm_drn_bwd_ws_bytes(N, B, k), dtype=torch.uint8, device=dev)
        self.out = torch.empty(B, 2, dtype=torch.float32, device=dev)
        self.net_out = torch.empty(B, 2, dtype=torch.float32, device=dev)
        self.dout = torch.empty(B, 2, dtype=torch.float32, device=dev)
        self._cap = (N, B)

    def train_step(self, batch: EventBatch, max_event_nodes: int | None = None) -> torch.Tensor:
        """One optimisation step on a device-resident batch (x [N,11], y [B,>=2], ptr [B+1] i32); returns the loss as a
        0-dim device tensor (no sync)."""
        x, y, p = batch.x, batch.y, batch.ptr
        N, B = x.shape[0], y.shape[0]
        men = max_event_nodes if max_event_nodes is not None else getattr(batch, "max_event_nodes", None)
        if men is None:
            men = int((p[1:] - p[:-1]).max().item())
        if not self.net.training:
            self.net.train()
        self._reserve(N, B)
        l, st, net = lib(), stream(self.device), self.net
        bn = net.edgeconv1.nn[4]
        # the workspaces are allocated at their capacity: the layout is a function of (N, B), so pass exact sizes
        fws = l.gm_drn_ws_bytes(N, B, int(net.k))
        bws = l.gm_drn_bwd_ws_bytes(N, B, int(net.k))
        check(l.gm_drn_forward(C.byref(self._P), ptr(x), x.stride(0), ptr(p), N, B, int(men), 1, PRECISION[net.edge_precision],
                               float(bn.momentum if bn.momentum is not None else 0.1), float(bn.eps), ptr(self.fwd_ws), fws,
                               ptr(self.out), None, None, st), "gm_drn_forward")
        check(l.gm_drn_met_loss(ptr(self.out), ptr(y), y.stride(0), B, ptr(self.net_out), ptr(self.loss), ptr(self.dout), st),
              "gm_drn_met_loss")
        check(l.gm_drn_backward(C.byref(self._P), C.byref(self._G), ptr(x), x.stride(0), ptr(p), N, B, ptr(self.fwd_ws), fws,
                                ptr(self.bwd_ws), bws, ptr(self.dout), st), "gm_drn_backward")
        self.step_count += 1
        g = self.optimizer.param_groups[0]
        check(l.gm_adamw_step(ptr(self.flat), ptr(self.flat_grad), ptr(self.exp_avg), ptr(self.exp_avg_sq), self.flat.numel(),
                              float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
                              float(g["weight_decay"]), self.step_count, 1.0, st), "gm_adamw_step")
        return self.loss

    def train_step_from_host(self, host_batch: EventBatch, max_event_nodes: int | None = None) -> float:
        dev = self.device
        x = host_batch.x.to(dev, non_blocking=True)
        y = host_batch.y.to(dev, non_blocking=True)
        p = host_batch.ptr.to(dev, non_blocking=True)
        men = max_event_nodes if max_event_nodes is not None else int((host_batch.ptr[1:] - host_batch.ptr[:-1]).max())
        loss = self.train_step(EventBatch(x, y, None, p), men)
        self._host_loss.copy_(loss, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return float(self._host_loss)
