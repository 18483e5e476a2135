"""HvqTrainer: the training step of the fork's live tokenizer (HierarchicalVQVAE forward + compute_vqvae_loss + backward +
AdamW; the loop body of robomimic/algo/icl.py:1437-1526 for the tokenizer alone) as ONE CUDA graph.

The eager step costs ~125 launches of liblipvq kernels plus ~200 small torch kernels and two device-to-host copies
(replace_dead_codes): at the reference's own batch (16 chunks of 10 steps) that is 7.8 ms of host time for under a
millisecond of GPU work.  The trainer runs the same module code with `_graph_safe` set (no host round trips: codebooks are
initialised before capture, dead codes are replaced with data-independent shapes -- see
HierarchicalVQVAE.replace_dead_codes_static for what that changes) and replays one captured graph per step.

    tr = HvqTrainer(model, lr=1e-3, weight_decay=1e-4, lambda_rec=0.002)
    loss = tr.step(actions)          # actions [B, T, A] on the model's device; returns the step's losses (device scalars)

Shapes are fixed by the first call; a call with another shape re-captures.  `graph=False` runs the same step eagerly (A/B
runs, debugging)."""
import torch

from . import _cabi as A  # noqa: F401  (importing it loads liblipvq.so or fails loudly)


class HvqTrainer:
    def __init__(self, model, lr=1e-3, weight_decay=1e-4, lambda_rec=1.0, graph=True, warmup=3):
        if next(model.parameters()).device.type != "cuda":
            raise RuntimeError("HvqTrainer needs the model on a CUDA device (no CPU path)")
        self.model = model
        self.lambda_rec = float(lambda_rec)
        self.opt = torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=weight_decay, capturable=bool(graph))
        self.use_graph = bool(graph)
        self.warmup = int(warmup)
        self._graph = None
        self._x = None
        self._losses = None

    def _eager_step(self, x):
        m = self.model
        self.opt.zero_grad(set_to_none=False)
        out = m(x, training=True)
        losses = m.compute_vqvae_loss(out, x, lambda_rec=self.lambda_rec)
        losses["vqvae_loss"].backward()
        self.opt.step()
        return {k: v.detach() for k, v in losses.items()}

    def _capture(self, x):
        m = self.model
        m.train()
        if not bool(m.codebooks_initialized):          # host-side, once, before anything is captured
            with torch.no_grad():
                m.initialize_codebooks_with_current_batch(m.encoder(x))
        m._graph_safe = True
        self._x = x.clone()
        side = torch.cuda.Stream(device=x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side):                   # warm-up on a side stream: allocator pools, optimizer state,
            for _ in range(self.warmup):                # kernel attributes and workspaces exist before capture
                self._eager_step(self._x)
        torch.cuda.current_stream(x.device).wait_stream(side)
        torch.cuda.synchronize(x.device)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._losses = self._eager_step(self._x)

    def step(self, x):
        if not self.use_graph:
            self.model._graph_safe = True
            if not bool(self.model.codebooks_initialized):
                with torch.no_grad():
                    self.model.initialize_codebooks_with_current_batch(self.model.encoder(x))
            return self._eager_step(x)
        if self._graph is None or self._x.shape != x.shape:
            self._capture(x)                            # (the warm-up steps of a capture are real training steps)
        self._x.copy_(x, non_blocking=True)
        self._graph.replay()
        return self._losses
