"""Lightning-free stand-in for the reference's `lightning_module.LightningModule` (lightning_module.py:23-203) that runs
the forward-only step bodies on the B200 module.

`pytorch_lightning` is not in this image and the optimizer / backward half of training is SURVEY 8(f) rank 4, so this
class keeps the reference's attribute and method names for the parts that are forward-only and drives them with a
20-line loop instead of a Trainer:

    reference                                         here
    ------------------------------------------------  -----------------------------------------------------------------
    __init__(config, exp_dir)          :24-38         same; model = transformertts_b200.TransformerTTS(**config['model']),
                                                      criterion = transformertts_b200.loss.TransformerTTSLoss(**config['loss'])
    forward(...)                       :41-42         same
    validation_step(batch, batch_idx)  :102-130       same body: prepare_batch -> forward -> criterion -> losses list
    on_validation_epoch_start/_end     :133-157       same: mean loss, then `model.inference(phoneme[:1], phoneme_lens[:1])`
    training_step(batch, batch_idx)    :45-86         `training_step_forward`: the no-grad teacher-forcing pass (:53-59),
                                                      the scheduled-sampling mix (:62-74), the second forward and the loss
                                                      (:77-78) -- without autograd.  DIFFERENCE, stated: the reference runs
                                                      both passes with the module in train mode (dropout on, BatchNorm batch
                                                      statistics); the B200 module is eval-only, so both passes here use
                                                      dropout-off / running statistics.
    plots (utils/plot.py, matplotlib)                 optional callables in `plots`; never imported (UI, out of scope)
    logger / self.log / lr_schedulers()               optional `log_fn(str)`; metrics kept in `self.logged`
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Iterable, Optional

import torch

from .api import TransformerTTS
from .data import PinnedFeeder
from .loss import TransformerTTSLoss
from .util import apply_teacher_forcing, get_teacher_forcing_ratio, prepare_batch


class LightningModule:
    def __init__(self, config: Dict[str, Any], exp_dir: Optional[str] = None, device="cuda",
                 plots: Optional[Dict[str, Callable]] = None, log_fn: Optional[Callable[[str], None]] = None):
        self.device = torch.device(device)
        self.model = TransformerTTS(**config["model"], device=self.device).eval()       # :30
        self.criterion = TransformerTTSLoss(**config["loss"]).to(self.device)           # :31
        self.config = config
        self.exp_dir = exp_dir
        self.train_losses, self.valid_losses = [], []                                   # :35-36
        self.example_batch = None
        self.log_interval = config["training"].get("log_interval", 100)                 # :38
        self.current_epoch = 0
        self.total_train_batches = self.total_valid_batches = 0
        self.plots = plots or {}
        self.log_fn = log_fn
        self.logged: Dict[str, float] = {}

    # ------------------------------------------------------------------ :41-42
    def forward(self, phoneme, melspec, phoneme_lens, melspec_lens):
        return self.model(phoneme, melspec, phoneme_lens, melspec_lens)

    __call__ = forward

    def _plot(self, name, *args):
        fn = self.plots.get(name)
        if fn is not None:
            fn(*args)

    # ------------------------------------------------------------------ :45-86 without the backward
    @torch.no_grad()
    def training_step_forward(self, batch, batch_idx, seed=None):
        phoneme, melspec, phoneme_lens, melspec_lens = prepare_batch(batch, self.device)
        pred_melspec = self.forward(phoneme=phoneme, melspec=melspec, phoneme_lens=phoneme_lens,
                                    melspec_lens=melspec_lens)["pred_melspec"]          # :53-59
        p_tf = get_teacher_forcing_ratio(epoch=self.current_epoch + 1,
                                         total_epochs=self.config["training"]["num_epochs"],
                                         mode=self.config["training"]["teacher_forcing_mode"], cycles=1)   # :62-67
        mel_mixed = apply_teacher_forcing(pred_melspec=pred_melspec, melspec=melspec, melspec_lens=melspec_lens,
                                          p_tf=p_tf, device=self.device, seed=seed)     # :68-74
        output = self.forward(phoneme, mel_mixed, phoneme_lens, melspec_lens)           # :77
        loss = self.criterion(output, melspec, melspec_lens)                            # :78
        if batch_idx % self.log_interval == 0:
            self._plot("plot_mels_scheduled", mel_mixed, melspec, self.current_epoch + 1, self.exp_dir)
            self.log_step_info(batch_idx, "TRAIN", loss)
        self.train_losses.append(loss["total"].item())                                  # :84
        return loss["total"], mel_mixed

    # ------------------------------------------------------------------ :102-130
    @torch.no_grad()
    def validation_step(self, batch, batch_idx):
        if self.example_batch is None:
            self.example_batch = batch
        phoneme, melspec, phoneme_lens, melspec_lens = prepare_batch(batch, self.device)
        output = self.forward(phoneme=phoneme, melspec=melspec, phoneme_lens=phoneme_lens, melspec_lens=melspec_lens)
        loss = self.criterion(output, melspec, melspec_lens)
        if batch_idx == 0:
            ep = self.current_epoch + 1
            self._plot("plot_mels_batch", output["pred_melspec"], melspec, ep, self.exp_dir)
            self._plot("plot_alignments_batch", output["alignments"], ep, self.exp_dir, 4)
            self._plot("plot_alignment_single", output["alignments"], 0, ep, self.exp_dir)
        if batch_idx % self.log_interval == 0:
            self.log_step_info(batch_idx, "VALID", loss)
        self.valid_losses.append(loss["total"].item())
        return loss["total"]

    def on_validation_epoch_start(self, total_valid_batches: int = 0):                  # :133-134
        self.total_valid_batches = total_valid_batches

    @torch.no_grad()
    def on_validation_epoch_end(self):                                                  # :137-157
        avg_valid_loss = sum(self.valid_losses) / len(self.valid_losses)
        self._log("[{:3d}] [{:04}/{:04}] | VALID | Average Loss: {:9.5f}".format(
            self.current_epoch + 1, self.total_valid_batches, self.total_valid_batches, avg_valid_loss))
        self.logged["val_loss"] = avg_valid_loss
        self.valid_losses.clear()
        pred_melspec = None
        if self.example_batch:
            phoneme, melspec, phoneme_lens, _ = prepare_batch(self.example_batch, self.device)
            pred_melspec = self.model.inference(phoneme[:1], phoneme_lens[:1])["pred_melspec"]          # :154
            self._plot("plot_mels_single", pred_melspec[0], melspec[0], self.current_epoch + 1, self.exp_dir)
        return avg_valid_loss, pred_melspec

    # ------------------------------------------------------------------ :181-203
    def _log(self, msg: str):
        if self.log_fn is not None:
            self.log_fn(msg)

    def log_step_info(self, batch_idx: int, phase: str, loss: Dict[str, torch.Tensor]):
        if phase not in ["TRAIN", "VALID"]:
            raise ValueError("phase must be either 'TRAIN' or 'VALID'")
        total = self.total_train_batches if phase == "TRAIN" else self.total_valid_batches
        if batch_idx % self.log_interval == 0 or batch_idx == total - 1:
            self._log("[{:3d}] [{:04}/{:04}] | {:>5} | total: {:9.5f} | mel: {:9.5f} | post: {:9.5f} | stop: {:9.5f}".format(
                self.current_epoch + 1, batch_idx, total, phase, loss["total"].item(), loss["pred_mel"].item(),
                loss["post_mel"].item(), loss["stop"].item()))

    # ------------------------------------------------------------------ the Trainer's validation loop, 1 device
    def run_validation(self, batches: Iterable[dict], mel_dtype: Optional[torch.dtype] = None):
        """What `Trainer.validate` does around the three hooks: every collated batch goes through the pinned
        double-buffered feeder (data.PinnedFeeder) and `validation_step`; returns (mean loss, example inference)."""
        batches = list(batches)
        self.on_validation_epoch_start(len(batches))
        for i, b in enumerate(PinnedFeeder(batches, self.device, mel_dtype=mel_dtype)):
            if self.example_batch is None:
                self.example_batch = batches[0]
            self.validation_step(b, i)
        return self.on_validation_epoch_end()
