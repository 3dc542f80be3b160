"""Step harness: the per-step training path of the three reference scripts, restated with only
packages that exist on this image (torch + transformers model / scheduler).

The literal scripts cannot run here (they import valohai, accelerate and evaluate, and
train-torchrun.py uses Trainer kwargs removed in transformers 5), so the
benchmark, the parity tests and the CPU baseline drive this harness instead. It follows, line
by line:

  semantics="accelerate"  train-accelerator.py:174-187 (two AdamW groups, both weight_decay 0,
                          lr 5e-5), :200-205 (linear schedule, 1 warm-up step), :217-226
                          (forward -> loss -> backward -> optimizer.step -> lr_scheduler.step
                          -> zero_grad), DDP wrap = what accelerator.prepare does (:189)
  semantics="task"        train-task.py:247-272 (same optimizer/schedule), :288-297 (forward,
                          loss.item() EVERY step, backward, average_gradients, optimizer.step,
                          lr_scheduler.step, zero_grad); no DDP wrapper (:65-69)
  semantics="trainer"     train-torchrun.py:115-128 + HF Trainer's step:
                          gradient accumulation with no_sync on all but the last micro-batch,
                          loss / GA, DDP mean, clip_grad_norm_(1.0), AdamW(weight_decay=0.01 on
                          non-LayerNorm / non-bias parameters), linear schedule, zero_grad

impl="b200" attaches the three plug-ins of this package through patch.py (CUDA only);
impl="stock" leaves torch / NCCL / ATen in place (the optimisation target on GPU and, on CPU
with gloo, the reference's own CPU path).
"""
from __future__ import annotations

import contextlib
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist

from . import patch
from .collate import Seq2SeqCollator
from .synthetic import build_model, make_features


@dataclass
class StepConfig:
    model: str = "flan-t5-large"
    semantics: str = "trainer"          # accelerate | task | trainer
    impl: str = "b200"                  # b200 | stock
    device: str = "cuda"
    batch_size: int = 8                 # per device, per micro-batch
    grad_accum: int = 1
    label_mode: str = "REF"
    bf16: bool = True                   # autocast with fp32 master parameters
    dropout: float = 0.1
    lr: float = 5e-5
    warmup_steps: int = 1
    total_steps: int = 1000
    max_grad_norm: float = 1.0          # trainer semantics only
    fused_adamw: Optional[bool] = None  # stock impl: torch's fused kernel (what Trainer selects)
    wire: str = "fp32"                  # b200 impl: dtype on NVLink
    bucket_cap_mb: Optional[float] = None  # DDP bucket size; None = the plug-in's / torch's default
    fused_lm_head: Optional[bool] = None   # b200 impl: lm_head + CE as one tcgen05 kernel (None: B2_FUSED_LM_HEAD)
    cuda_graph: Optional[bool] = None      # b200 impl: replay forward + backward from CUDA graphs (None: B2_CUDA_GRAPH)


def _no_decay(name: str) -> bool:
    # Trainer.get_decay_parameter_names: LayerNorm-class parameters and biases are not decayed
    return "layer_norm" in name or "bias" in name


class StepHarness:
    def __init__(self, cfg: StepConfig):
        self.cfg = cfg
        self.device = torch.device(cfg.device)
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        if cfg.impl == "b200":
            if self.device.type != "cuda":
                raise RuntimeError("impl='b200' has no CPU path")
            # the same attach path an unmodified script gets from launch.py: CrossEntropyLoss,
            # torch.optim.AdamW, clip_grad_norm_ and DistributedDataParallel.__init__ are rebound
            patch.remove()
            patch.apply(comm_kwargs={"wire": cfg.wire}, fused_lm_head=cfg.fused_lm_head, cuda_graph=cfg.cuda_graph)
        elif cfg.impl == "stock":
            patch.remove()
        else:
            raise ValueError(cfg.impl)
        from transformers import get_scheduler

        self.model = build_model(cfg.model, cfg.dropout).to(self.device)
        self.model.train()
        named = list(self.model.named_parameters())
        if cfg.semantics in ("accelerate", "task"):
            no_decay = ["bias", "LayerNorm.weight"]  # the scripts' own lists; both groups get 0.0
            groups = [
                {"params": [p for n, p in named if not any(nd in n for nd in no_decay)], "weight_decay": 0.0},
                {"params": [p for n, p in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0},
            ]
            groups = [g for g in groups if g["params"]]
            kw = {}
        elif cfg.semantics == "trainer":
            groups = [
                {"params": [p for n, p in named if not _no_decay(n)], "weight_decay": 0.01},
                {"params": [p for n, p in named if _no_decay(n)], "weight_decay": 0.0},
            ]
            kw = {"betas": (0.9, 0.999), "eps": 1e-8}
        else:
            raise ValueError(cfg.semantics)
        if cfg.impl == "b200":
            # what the scripts' ``from torch.optim import AdamW`` resolves to after patch.apply()
            self.optimizer = torch.optim.AdamW(groups, lr=cfg.lr, **kw)
        else:
            fused = cfg.fused_adamw
            if fused is None:
                fused = cfg.semantics == "trainer" and self.device.type == "cuda"
            self.optimizer = torch.optim.AdamW(groups, lr=cfg.lr, fused=fused or None, **kw)
        self.scheduler = get_scheduler("linear", self.optimizer, num_warmup_steps=cfg.warmup_steps,
                                       num_training_steps=cfg.total_steps)
        self.ddp = None
        self.hook_state = None
        self.comm = None
        if self.world > 1:
            if cfg.semantics == "task":
                if cfg.impl == "b200":
                    from .comm import P2PComm

                    self.comm = P2PComm(wire=cfg.wire)
            else:
                from torch.nn.parallel import DistributedDataParallel as DDP

                ids = [self.device.index] if self.device.type == "cuda" else None
                extra = {} if cfg.bucket_cap_mb is None else {"bucket_cap_mb": cfg.bucket_cap_mb}
                self.ddp = DDP(self.model, device_ids=ids, find_unused_parameters=False, **extra)
                if cfg.impl == "b200":
                    # registered by the patched DDP.__init__ (patch.patch_ddp), not by hand
                    self.hook_state = self.ddp._b200_hook_state
        self.collate = Seq2SeqCollator(pad_to_multiple_of=8 if cfg.label_mode == "MASK" else None,
                                       pin_memory=self.device.type == "cuda")
        self.steps_done = 0

    # -- data ----------------------------------------------------------------------------
    def host_batches(self, n_micro: int, seed: int = 1234) -> list[dict]:
        """n_micro collated micro-batches in (pinned) host memory, seeded 1234 + rank."""
        feats = make_features(n_micro * self.cfg.batch_size, rank=self.rank, label_mode=self.cfg.label_mode,
                              seed=seed)
        b = self.cfg.batch_size
        return [self.collate(feats[i * b:(i + 1) * b]) for i in range(n_micro)]

    def to_device(self, batch: dict) -> dict:
        return {k: v.to(self.device, non_blocking=True) for k, v in batch.items()}

    @staticmethod
    def batch_bytes(batch: dict) -> int:
        return sum(v.numel() * v.element_size() for v in batch.values())

    # -- one optimizer step ----------------------------------------------------------------
    def _forward(self, batch: dict) -> torch.Tensor:
        net = self.ddp if self.ddp is not None else self.model
        ctx = (torch.autocast(self.device.type, dtype=torch.bfloat16) if self.cfg.bf16
               else contextlib.nullcontext())
        with ctx:
            return net(**batch).loss

    def step(self, micro_batches: list[dict]) -> torch.Tensor:
        """One optimizer step over ``grad_accum`` micro-batches (already on the device or on the
        host: host batches are copied here, inside whatever region the caller is timing)."""
        cfg = self.cfg
        assert len(micro_batches) == cfg.grad_accum
        loss_out = None
        for i, mb in enumerate(micro_batches):
            if next(iter(mb.values())).device != self.device:
                mb = self.to_device(mb)
            last = i == cfg.grad_accum - 1
            sync_ctx = (self.ddp.no_sync() if (self.ddp is not None and not last) else contextlib.nullcontext())
            with sync_ctx:
                loss = self._forward(mb)
                if cfg.semantics == "task":
                    loss.item()  # train-task.py:291: host sync every step
                (loss / cfg.grad_accum if cfg.grad_accum > 1 else loss).backward()
            loss_out = loss.detach() if loss_out is None else loss_out + loss.detach()
        if cfg.semantics == "task" and self.world > 1:
            if cfg.impl == "b200":
                from .comm import average_gradients

                average_gradients(self.model, self.comm)
            else:
                size = float(self.world)  # train-task.py:65-69 verbatim semantics
                for p in self.model.parameters():
                    dist.all_reduce(p.grad.data, op=dist.ReduceOp.SUM)
                    p.grad.data /= size
        if cfg.semantics == "trainer" and cfg.max_grad_norm:
            # what HF Trainer calls (trainer.py:1753-1757); with impl="b200" the name is rebound by
            # patch.patch_clip_grad_norm and routed to AdamW.clip_grad_norm_ (norm from the
            # all-reduce's fold or one read-only pass; the multiply happens inside the AdamW launch)
            torch.nn.utils.clip_grad_norm_(self.model.parameters(), cfg.max_grad_norm)
        self.optimizer.step()
        self.scheduler.step()
        self.optimizer.zero_grad()
        self.steps_done += 1
        return loss_out / cfg.grad_accum

    def close(self) -> None:
        if self.hook_state is not None:
            self.hook_state.comm.close()
        if self.comm is not None:
            self.comm.close()
