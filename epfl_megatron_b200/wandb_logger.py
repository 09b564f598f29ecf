"""TensorBoard-API shim that buffers scalars per step and emits one ``wandb.log`` per step
(megatron/wandb_logger.py:13-160 behaviour). ``wandb`` is imported lazily; without it (or
without an API key) scalars are kept in memory and optionally mirrored to TensorBoard."""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Optional


@dataclass
class WandBConfig:
    project: Optional[str] = None
    entity: Optional[str] = None
    run_id: Optional[str] = None
    resume: bool = False
    api_key: Optional[str] = None
    tensorboard_dir: Optional[str] = None
    config: dict = field(default_factory=dict)

    @staticmethod
    def from_args(args):
        cfg = {k: (str(v) if not isinstance(v, (int, float, str, bool, type(None))) else v)
               for k, v in vars(args).items()}
        return WandBConfig(project=args.wandb_project, entity=args.wandb_entity, run_id=args.wandb_id,
                           resume=args.wandb_resume, api_key=args.wandb_api_key,
                           tensorboard_dir=getattr(args, "tensorboard_dir", None), config=cfg)


class WandbTBShim:
    def __init__(self, config: WandBConfig):
        self.cfg = config
        self._step = None
        self._buffer = {}
        self.history = []
        self._wandb = None
        self._tb = None
        key = config.api_key or os.environ.get("WANDB_API_KEY")
        try:
            import wandb  # noqa: WPS433
            if key:
                wandb.login(key=key)
                wandb.init(project=config.project, entity=config.entity, id=config.run_id,
                           resume="must" if config.resume else None, config=config.config)
                self._wandb = wandb
        except Exception as e:  # noqa: BLE001
            print(f"WARNING: wandb unavailable ({type(e).__name__}); scalars are buffered locally", flush=True)
        if config.tensorboard_dir:
            try:
                from torch.utils.tensorboard import SummaryWriter
                self._tb = SummaryWriter(log_dir=config.tensorboard_dir)
            except Exception:  # noqa: BLE001
                self._tb = None

    def add_scalar(self, name, value, step=None, *a, **kw):
        if self._tb is not None:
            self._tb.add_scalar(name, value, step)
        if " vs " in name:   # the "... vs samples" twins are dropped
            return
        if self._step is not None and step != self._step:
            self.flush_all()
        self._step = step
        try:
            self._buffer[name] = float(value)
        except (TypeError, ValueError):
            pass

    def add_text(self, name, value, global_step=None):
        if self._tb is not None:
            self._tb.add_text(name, value, global_step)

    def flush_all(self):
        if not self._buffer:
            return
        self.history.append((self._step, dict(self._buffer)))
        if self._wandb is not None:
            self._wandb.log(self._buffer, step=self._step)
        self._buffer = {}

    def flush(self):
        self.flush_all()
