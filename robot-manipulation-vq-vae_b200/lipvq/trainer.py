"""Tokenizer training step as the reference's algo layer runs it, expressed as one launch sequence.

Reference protocol (robomimic/algo/icl.py:1437-1441, 1466, 1520-1522): a dedicated
``AdamW(vq_vae_model.parameters(), lr=1e-3, weight_decay=1e-4)``; per step zero_grad -> forward ->
``_vq_vae_loss.backward()`` -> ``vq_optimizer.step()``; no gradient clipping on the tokenizer.
With an EMA quantizer (LFQQuantizerEMA, robomimic/models/vq_vae/hq_vae.py:224-281) the codebook is
overwritten from the EMA buffers each step and still takes its AdamW update, as the reference does.

Layout in HBM: parameters, gradients and both Adam moments are single flat fp32 slabs (the module's
nn.Parameters are re-pointed at views of the parameter slab, so ``state_dict`` keeps working).  The
reduction slab is ``[grads | counts(K) | sums(K*D) | losses(4)]`` (parallel.SlabLayout) so a
data-parallel step needs exactly ONE ``all_reduce(sum)`` (NCCL over NVLink); every rank then applies
the identical EMA + AdamW update and replicas stay bit-identical.  The two halves of the step (before /
after the reduction) are captured as CUDA graphs.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import _cabi as A
from . import ops, t1
from .modules import LFQQuantizerEMA, LLFQVAE_V4, _T1_KEYS
from .parallel import PeerSlab, SlabLayout, all_reduce_slab


class TokenizerTrainer:
    def __init__(self, model: LLFQVAE_V4, batch_tokens: int, lr: float = 1e-3, weight_decay: float = 1e-4,
                 betas=(0.9, 0.999), eps: float = 1e-8, use_graph: bool = True,
                 process_group: Optional[dist.ProcessGroup] = None, search_algo: int = A.ALGO_AUTO):
        self.model = model
        dev = next(model.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("TokenizerTrainer needs the model on a CUDA device (no CPU path)")
        self.device = dev
        self.N = int(batch_tokens)
        F_, D, _, H = model._dims
        K = model.quantizer.codebook.shape[0]
        self.F, self.D, self.K, self.H = F_, D, K, H
        self.ema = isinstance(model.quantizer, LFQQuantizerEMA)
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        self.search_algo = search_algo
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1

        named = dict(model.named_parameters())
        self.layout = SlabLayout(list(_T1_KEYS), [tuple(named[k].shape) for k in _T1_KEYS], K, D, self.ema)
        L = self.layout
        self.flat_p = torch.empty(L.P, dtype=torch.float32, device=dev)
        # reduction slab: phase A accumulates into ``red``; phase B reads the cross-rank sum from ``red_out``.  With the
        # NVLink peer all-reduce (default for world > 1) red lives in symmetric memory and red_out is a local buffer;
        # with NCCL (fallback, or LIPVQ_DP_COLLECTIVE=nccl) the all-reduce is in place and both are the same tensor.
        self.peer = None
        self.collective = "none"
        if self.world > 1:
            # LIPVQ_DP_COLLECTIVE: "peer" (default: our one-shot NVLink kernel; NCCL only if symmetric memory cannot be
            # set up on EVERY rank, and then loudly), "peer!" (no fallback: raise), "nccl"
            mode = os.environ.get("LIPVQ_DP_COLLECTIVE", "peer")
            err = None
            if mode in ("peer", "peer!"):
                try:
                    self.peer = PeerSlab(L.length, dev, process_group)
                except Exception as e:  # noqa: BLE001 -- any failure of the symmetric-memory plumbing
                    err = e
                # the ranks must agree: a subset spinning in the peer kernel while the rest sit in NCCL is a hang
                ok = torch.tensor([0 if self.peer is None else 1], dtype=torch.int32, device=dev)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=process_group)
                if int(ok.item()) == 0:
                    self.peer = None
                    msg = f"lipvq: NVLink peer all-reduce unavailable on at least one rank ({err}); "
                    if mode == "peer!":
                        raise RuntimeError(msg + "LIPVQ_DP_COLLECTIVE=peer! forbids the NCCL fallback")
                    import warnings
                    warnings.warn(msg + "ALL ranks use NCCL for the data-parallel reduction")
            elif mode != "nccl":
                raise ValueError(f"LIPVQ_DP_COLLECTIVE={mode!r}: expected peer, peer! or nccl")
            self.collective = "peer" if self.peer is not None else "nccl"
        self.red = self.peer.slab if self.peer is not None else torch.zeros(L.length, dtype=torch.float32, device=dev)
        self.red_out = self.peer.out if self.peer is not None else self.red
        self.flat_g = self.red_out[:L.P]
        self.flat_m = torch.zeros(L.P, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(L.P, dtype=torch.float32, device=dev)
        self.p: Dict[str, torch.Tensor] = {}
        self.g: Dict[str, torch.Tensor] = {}
        with torch.no_grad():
            for k in _T1_KEYS:
                view = L.grad_view(self.flat_p, k)
                view.copy_(named[k].detach())
                named[k].data = view  # the module now lives in the slab
                self.p[k] = view
                self.g[k] = L.grad_view(self.red, k)
        if self.ema:
            self.counts, self.sums = L.counts(self.red), L.sums(self.red)
            self.counts_out, self.sums_out = L.counts(self.red_out), L.sums(self.red_out)
        self.losses_in = L.losses(self.red)
        self.losses_red = L.losses(self.red_out)
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.scratch = torch.zeros(1, dtype=torch.float64, device=dev)
        self.bufs = t1.T1Buffers(self.N, F_, D, K, H, dev)
        self.bufs.alloc_backward()
        # buffers that must be zero when a step starts (sum-of-squares accumulators, the Lipschitz layer's dense
        # gradient) share one arena so one fill clears them; the loss scalars are written straight into the slab
        arena = torch.zeros(D * H + 4, dtype=torch.float32, device=dev)
        self._arena = arena
        self.bufs.t["G"] = arena[:D * H].view(D, H)
        self.bufs.t["sq"] = arena[D * H:D * H + 4].view(torch.float64)
        self.bufs.zeroed_by_caller = True
        self.bufs.t["losses"] = self.losses_in
        self.x_static = torch.empty(self.N, F_, dtype=torch.float32, device=dev)
        # The codebook changes once per step (EMA overwrite + AdamW, phase B): its search-side preparation runs right
        # there, centred on that step's latents, and the next step's search starts from the prepared image.  Anything
        # that changes the parameters behind the trainer's back (load_state_dict, manual edits) must call refresh().
        self.search_prepared = (self.bufs.search_prepared_ok and search_algo in (A.ALGO_AUTO, A.ALGO_TCGEN05)
                                and os.environ.get("LIPVQ_TRAINER_PREPARED", "1") != "0")
        # (the centre of the prepared contraction is a batch of latents: until the first step has produced one, the step
        #  uses the self-contained search, which takes its centre from the rows it is given)
        self._centre_ok = False
        self.use_graph = use_graph
        self._graph_a = None  # zero grads, forward, EMA statistics, backward
        self._graph_b = None  # EMA apply + AdamW
        self._graph_full = None  # world > 1: phase a + all-reduce + phase b in one graph
        self._graph_tok = None
        self._graph_tok_fast = None
        self._graph_tok_bf16 = None
        self._graph_tok_fast_bf16 = None
        self.graph_collective = os.environ.get("LIPVQ_GRAPH_COLLECTIVE", "1") != "0"
        # input prefetch (step(x, prefetch=x_next)): the next batch's H2D copy runs on its own stream under this step
        self._stage = None
        self._copy_stream = None
        self._stage_ready = None
        self._stage_free = None
        self._stage_src = None
        self.steps_done = 0

    # -- the launch sequence -------------------------------------------------------------------
    def _phase_a(self):
        if self.peer is not None:
            self.peer.clear()  # (waits for the peers' "slab read" signals of the last reduction, then zeroes)
        else:
            self.red.zero_()
        self._arena.zero_()
        t1.forward(self.p, self.x_static, self.bufs, algo=self.search_algo,
                   search_prepared=self.search_prepared and self._centre_ok)
        if self.ema:
            t1.ema_stats(self.bufs, self.counts, self.sums)
        t1.backward(self.p, self.g, self.x_static, self.bufs)

    def _reduce(self):
        if self.peer is not None:
            self.peer.all_reduce()
        else:
            all_reduce_slab(self.red, self.pg)

    def _phase_b(self):
        if self.ema:
            q = self.model.quantizer
            ops.ema_apply(self.counts_out, self.sums_out, q.ema_cluster_size, q.ema_codebook, self.p["quantizer.codebook"],
                          q.decay, q.epsilon, A.EMA_SIMPLE, self.scratch)
        ops.counter_add(self.step_dev, 1)
        ops.adamw_step(self.flat_p, self.flat_g, self.flat_m, self.flat_v, 0, self.step_dev, self.lr, self.betas,
                       self.eps, self.wd, 1.0 / self.world)
        if self.search_prepared:  # the codebook has its final value of this step
            t1.prepare_search(self.p, self.bufs, center=self.bufs.t["z_e"])
            self._centre_ok = True

    def refresh(self) -> None:
        """Call after changing parameters outside step() (load_state_dict, ...): re-derives what is cached per
        parameter version (the prepared codebook of the search)."""
        self._tok_ready = False
        with torch.cuda.device(self.device):
            if self.search_prepared and self._centre_ok:  # (the latents of the last batch are still a good centre)
                t1.prepare_search(self.p, self.bufs, center=self.bufs.t["z_e"])

    def _snapshot(self):
        st = [self.flat_p.clone(), self.flat_m.clone(), self.flat_v.clone(), self.step_dev.clone()]
        if self.ema:
            q = self.model.quantizer
            st += [q.ema_cluster_size.clone(), q.ema_codebook.clone()]
        return st

    def _restore(self, st):
        self.flat_p.copy_(st[0]); self.flat_m.copy_(st[1]); self.flat_v.copy_(st[2]); self.step_dev.copy_(st[3])
        if self.ema:
            q = self.model.quantizer
            q.ema_cluster_size.copy_(st[4]); q.ema_codebook.copy_(st[5])
        if self.search_prepared and self._centre_ok:
            t1.prepare_search(self.p, self.bufs, center=self.bufs.t["z_e"])

    def _capture(self):
        # one eager pass on a side stream first (lazy module loading, allocator), then capture both halves;
        # the eager pass advanced the optimiser state, so put it back afterwards
        st = self._snapshot()
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            self._phase_a()
            if self.world > 1:
                self._reduce()  # also brings the NCCL communicator up before any capture
            self._phase_b()
        torch.cuda.current_stream(self.device).wait_stream(s)
        self._graph_full = None
        if self.world > 1 and self.graph_collective:
            # data-parallel: the whole step -- both halves AND the NCCL all-reduce between them -- is one graph, so a
            # step is a single launch with no host round-trip around the collective
            try:
                gf = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gf):
                    self._phase_a()
                    self._reduce()
                    self._phase_b()
                self._graph_full = gf
            except Exception as e:  # NCCL build without capture support: fall back to graph / eager collective / graph
                import warnings
                warnings.warn(f"lipvq: capturing the all-reduce failed ({e}); using two graphs around an eager collective")
                torch.cuda.synchronize()
        ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        if self._graph_full is None:
            with torch.cuda.graph(ga):
                self._phase_a()
            with torch.cuda.graph(gb):
                self._phase_b()
        self._restore(st)
        self._graph_a, self._graph_b = ga, gb

    def _load_input(self, x: torch.Tensor, prefetch: Optional[torch.Tensor]) -> None:
        cur = torch.cuda.current_stream(self.device)
        if self._stage_src is not None and x is self._stage_src:
            cur.wait_event(self._stage_ready)          # the copy issued during the previous step
            self.x_static.copy_(self._stage, non_blocking=True)
        else:
            self.x_static.copy_(x, non_blocking=True)
        self._stage_src = None
        if prefetch is not None:
            if tuple(prefetch.shape) != tuple(self.x_static.shape):
                raise RuntimeError(f"prefetch: expected shape {tuple(self.x_static.shape)}, got {tuple(prefetch.shape)}")
            if self._stage is None:
                self._stage = torch.empty_like(self.x_static)
                self._copy_stream = torch.cuda.Stream(device=self.device)
                self._stage_ready, self._stage_free = torch.cuda.Event(), torch.cuda.Event()
            self._stage_free.record(cur)                # staging buffer is free once x_static has been filled from it
            self._copy_stream.wait_event(self._stage_free)
            with torch.cuda.stream(self._copy_stream):
                self._stage.copy_(prefetch, non_blocking=True)
                self._stage_ready.record(self._copy_stream)
            self._stage_src = prefetch

    def step(self, x: torch.Tensor, prefetch: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One optimisation step on x [N,F] (device tensor, or pinned host tensor -> async H2D).
        ``prefetch``: the batch of the NEXT step (same shape); its host->device copy is started on a separate stream
        so it runs under this step's kernels -- pass the same tensor object as ``x`` to the next call.
        Returns a VIEW of the 4 loss scalars (loss, recon, commitment, codebook; summed over ranks -- divide
        by ``world``) that the next step overwrites."""
        if tuple(x.shape) != tuple(self.x_static.shape):
            raise RuntimeError(f"expected input of shape {tuple(self.x_static.shape)}, got {tuple(x.shape)}")
        with torch.cuda.device(self.device):  # the trainer's streams, graphs and kernels live on the model's device
            return self._step(x, prefetch)

    def _step(self, x: torch.Tensor, prefetch: Optional[torch.Tensor]) -> torch.Tensor:
        self._tok_ready = False  # the step updates the parameters after it clamped to_latent.W
        self._load_input(x, prefetch)
        if self.use_graph and self._graph_a is None:
            self._capture()
        if self.use_graph and self._graph_full is not None:
            self._graph_full.replay()
            self.steps_done += 1
            return self.losses_red
        if self.use_graph:
            self._graph_a.replay()
        else:
            self._phase_a()
        if self.world > 1:
            self._reduce()
        if self.use_graph:
            self._graph_b.replay()
        else:
            self._phase_b()
        self.steps_done += 1
        return self.losses_red

    @property
    def indices(self) -> torch.Tensor:
        return self.bufs.t["idx"]

    @torch.no_grad()
    def tokenize(self, x: torch.Tensor, prepared: bool = False, bf16: bool = False) -> torch.Tensor:
        """encode + VQ only on x [N,F] (bulk tokenisation) -> int64 [N] (a buffer the next call overwrites).
        ``prepared=True``: the parameters have not changed since the last step() / tokenize() / refresh() of this
        trainer -- the clamped to_latent weight and the prepared codebook of that call are reused (bulk tokenisation
        of a dataset with fixed weights: one encoder launch + one search launch per call).  ``bf16=True``: the encoder's
        matrix products take bf16 operands (fp32 accumulation; latents within 1e-2 of the fp32 path, so a token can
        differ where two codes are about equally near); the search on those latents is exact as always."""
        if tuple(x.shape) != tuple(self.x_static.shape):
            raise RuntimeError(f"expected input of shape {tuple(self.x_static.shape)}, got {tuple(x.shape)}")
        with torch.cuda.device(self.device):
            return self._tokenize(x, prepared and self.search_prepared and self._tok_ready, bool(bf16))

    _tok_ready = False  # Wn and the prepared codebook in bufs are those of the current parameters

    def _tok_forward(self, fast: bool, bf16: bool = False) -> None:
        if fast:
            t1.forward(self.p, self.x_static, self.bufs, algo=self.search_algo, encode_only=True, search_prepared=True,
                       weights_prepared=True, encoder_bf16=bf16)
            return
        t1.forward(self.p, self.x_static, self.bufs, algo=self.search_algo, encode_only=True,
                   prepare_search_here=self.search_prepared, encoder_bf16=bf16)

    def _tokenize(self, x: torch.Tensor, fast: bool, bf16: bool = False) -> torch.Tensor:
        self.x_static.copy_(x, non_blocking=True)
        if not self.use_graph:
            self._tok_forward(fast, bf16)
            self._tok_ready = True
            return self.bufs.t["idx"]
        attr = ("_graph_tok_fast" if fast else "_graph_tok") + ("_bf16" if bf16 else "")
        if getattr(self, attr) is None:
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(s):
                self._tok_forward(fast, bf16)
            torch.cuda.current_stream(self.device).wait_stream(s)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._tok_forward(fast, bf16)
            setattr(self, attr, g)
        getattr(self, attr).replay()
        self._tok_ready = True
        return self.bufs.t["idx"]
