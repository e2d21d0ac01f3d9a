"""Training / evaluation loops with the reference Trainer's fields and methods
(modules/model/trainer/trainer.py:48-403): `train(after_epoch_funcs)`, `test(epoch_i, callbacks=)`,
`save_state_dict(path)`, `load_state_dict(path)`.

What differs underneath (same arithmetic):
  * the micro-batch body (trainer.py:276-280) is ONE fused call: forward, loss and backward
    kernels on the current stream, no autograd graph, no per-key `.item()`;
  * data parallelism is a bucketed NCCL all-reduce of the flat gradient buffer on a side stream,
    overlapped with the backward of the remaining layers, issued once per optimizer step
    (the reference's DDP wrapper, trainer.py:136-142, reduces on every micro-batch);
  * clip + AdamW + zero_grad (trainer.py:285-287) are two kernel passes;
  * loss scalars reach the host through one pinned async copy per micro-batch, read one
    iteration late so that the loop never blocks on the GPU;
  * `apex_level` / `sync_bn` are accepted and ignored: compute is bf16 with fp32 master weights
    and BERT has no BatchNorm.
"""
import functools
import logging
import os
import shutil
import time
from collections import defaultdict
from dataclasses import dataclass
from typing import Any, Optional

import torch
import torch.nn as nn
from torch.utils.data import DataLoader, DistributedSampler, RandomSampler, WeightedRandomSampler

from ..dataset import expand_compact_inputs
from .callback import TestCallback
from .meters import AverageMeter
from .reducer import DEFAULT_BUCKET_MB, GradReducer

logger = logging.getLogger(__file__)

_LOSS_KEYS = ('start_class', 'end_class', 'start_reg', 'end_reg', 'cls', 'loss')


def time_profiler(fun):
    @functools.wraps(fun)
    def wrapped(*args, **kwargs):
        t0 = time.perf_counter()
        try:
            return fun(*args, **kwargs)
        finally:
            logger.info(f'Execution of {fun.__name__} took {time.perf_counter() - t0:.3f} sec.')
    return wrapped


def linear_warmup_lambda(num_warmup_steps, num_training_steps):
    """lr multiplier of transformers.get_linear_schedule_with_warmup (trainer.py:125)."""
    def fn(step):
        if step < num_warmup_steps:
            return float(step) / float(max(1, num_warmup_steps))
        return max(0.0, float(num_training_steps - step) / float(max(1, num_training_steps - num_warmup_steps)))
    return fn


class _LossReadback:
    """Six loss scalars per micro-batch: device -> pinned host, asynchronously, double-buffered."""

    def __init__(self, device):
        self.host = [torch.zeros(6, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.event = [torch.cuda.Event() for _ in range(2)]
        self.i = 0

    def push(self, losses_dev):
        k = self.i & 1
        self.host[k].copy_(losses_dev, non_blocking=True)
        self.event[k].record()
        self.i += 1

    def pop_previous(self):
        """Values of the micro-batch before the last pushed one (None if there is none)."""
        if self.i < 2:
            return None
        k = (self.i - 2) & 1
        self.event[k].synchronize()
        return self.host[k].tolist()

    def pop_last(self):
        if self.i == 0:
            return None
        k = (self.i - 1) & 1
        self.event[k].synchronize()
        return self.host[k].tolist()


@dataclass
class Trainer:
    model: nn.Module
    loss: Any
    collate_fun: Any

    optimizer: Any = None

    train_dataset: Any = None
    test_dataset: Any = None

    writer_dir: Any = None

    device: Any = torch.device('cuda')

    local_rank: int = -1
    gpu_id: Optional[int] = None
    sync_bn: bool = False

    n_epochs: int = 0

    train_batch_size: int = 32
    test_batch_size: int = 32

    batch_split: int = 1
    n_jobs: int = 4

    warmup_coef: float = 0.01
    max_grad_norm: float = 1

    apex_level: str = None
    apex_verbosity: int = 1
    apex_loss_scale: float = None

    train_weights: defaultdict = None

    drop_optimizer: bool = False
    debug: bool = False

    def __post_init__(self):
        self.device = torch.device(self.device)
        if self.device.type != 'cuda':
            raise RuntimeError('this Trainer drives the sm_100a kernels: device must be a CUDA device '
                               '(the CPU path of the recipe is the reference itself)')
        if self.device.index is None:
            self.device = torch.device('cuda', torch.cuda.current_device())
        self.model = self.model.to(self.device)
        self.loss = self.loss.to(self.device)

        self.train_dataloader = self._init_dataloader(self.train_dataset, 'Train',
                                                      batch_size=int(self.train_batch_size // self.batch_split),
                                                      sampler=self._init_train_sampler(), drop_last=True)
        self.test_dataloader = self._init_dataloader(self.test_dataset, 'Test', batch_size=self.test_batch_size,
                                                     sampler=None, drop_last=False)

        self.scheduler = None
        if self.train_dataloader is not None and self.optimizer is not None and self.warmup_coef > 0:
            steps = self.n_epochs * len(self.train_dataloader) // self.batch_split
            warmup = int(steps * self.warmup_coef)
            logger.info(f'Warmup schedule is used. #Training steps: {steps}. #Warmup steps: {warmup}.')
            self.scheduler = torch.optim.lr_scheduler.LambdaLR(self.optimizer, linear_warmup_lambda(warmup, steps))
        logger.info(f'apex_level={self.apex_level} is accepted for compatibility; compute is bf16 with fp32 master '
                    f'weights, no loss scaling.')

        # The backward of an optimizer step's last micro-batch runs bucket by bucket: as a bucket of the flat gradient
        # becomes final it is all-reduced (N > 1 ranks) and its share of the global-norm sum of squares is taken, both
        # on a side stream under the backward of the remaining layers.
        self.reducer = None
        self.distributed = (self.local_rank != -1 and torch.distributed.is_initialized()
                            and torch.distributed.get_world_size() > 1)
        if self.optimizer is not None:
            eng = self.model.engine
            begins = [eng.phase_grad_begin(k) for k in range(eng.num_phases)]
            # bucket size: few large buckets when they are exchanged (reducer.DEFAULT_BUCKET_MB); a single rank only
            # overlaps the sum of squares with the backward, where one bucket per layer is best
            # exchange dtype: the reference's DDP moves fp32 gradients under apex O0 / O1 and 16-bit ones under O2 / O3
            # (trainer.py:128-142); BQA_GRAD_EXCHANGE=bf16|fp32 overrides
            exch = os.environ.get('BQA_GRAD_EXCHANGE', 'bf16' if self.apex_level in ('O2', 'O3') else 'fp32')
            self.reducer = GradReducer(eng.grads, begins, world=None if self.distributed else 1,
                                       exchange_dtype=torch.bfloat16 if exch == 'bf16' else torch.float32,
                                       min_bucket_mb=float(os.environ.get('BQA_BUCKET_MB', DEFAULT_BUCKET_MB if self.distributed else 25.0)))
            if self.distributed:
                # replicas must start identical (DDP broadcasts rank 0's parameters at wrap time)
                torch.distributed.broadcast(eng.params, src=0)
                eng.sync_weights()

        self.global_step = 0
        self.writer = self._init_writer(self.local_rank, self.writer_dir)
        self._loss_scale_cfg = None
        if self.debug:
            self.n_epochs = 2

    # ---- construction helpers ---------------------------------------------------------------
    def _init_train_sampler(self):
        if self.train_dataset is None:
            return None
        if self.local_rank == -1:
            w = None if self.train_weights is None else self.train_weights['sampler_weights']
            if w is None:
                sampler = RandomSampler(self.train_dataset)
            else:
                assert len(w) == len(self.train_dataset)
                sampler = WeightedRandomSampler(w, len(self.train_dataset))
        else:
            sampler = DistributedSampler(self.train_dataset)
        logger.info(f'Used train sampler: {type(sampler).__name__}.')
        return sampler

    def _init_dataloader(self, dataset, name, *, batch_size, sampler, drop_last):
        if dataset is None:
            return None
        logger.info(f'{name} dataset len: {len(dataset)}. #JOBS: {self.n_jobs}.')
        return DataLoader(dataset, batch_size=batch_size, num_workers=self.n_jobs, sampler=sampler,
                          drop_last=drop_last, shuffle=False, collate_fn=self.collate_fun, pin_memory=True)

    @staticmethod
    def _init_writer(local_rank, writer_dir):
        if writer_dir is None or local_rank not in [-1, 0]:
            return None
        try:
            from torch.utils.tensorboard import SummaryWriter
        except Exception as e:  # tensorboard is optional
            logger.warning(f'TensorBoard writer is unavailable: {e}')
            return None
        logger.warning(f'Directory {writer_dir} will be cleaned before SummaryWriter initialization.')
        shutil.rmtree(writer_dir, ignore_errors=True)
        return SummaryWriter(log_dir=writer_dir)

    # ---- small helpers ------------------------------------------------------------------------
    def _get_lr(self):
        return self.optimizer.param_groups[0]['lr']

    @staticmethod
    def _get_console_str(losses):
        return ', '.join(f'{k}: {v() if isinstance(v, AverageMeter) else v:.3e}' for k, v in losses.items())

    def _update_writer(self, losses, *, prefix=None):
        if self.writer is not None:
            for k, v in losses.items():
                self.writer.add_scalar(f'{prefix}/{k}', v() if isinstance(v, AverageMeter) else v,
                                       global_step=self.global_step)

    def _clip_grad_norm(self):
        if hasattr(self.optimizer, 'prepare_clip'):
            self.optimizer.prepare_clip(self.max_grad_norm)
        else:
            torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.max_grad_norm)

    def set_train(self):
        # trainer.py:226-233: in fine-tuning mode only the selected sub-modules go to train mode
        trainable = getattr(self.model, 'list_of_trainable_modules', None)
        if self.apex_level is None and trainable is not None:
            assert isinstance(trainable, list) and all(isinstance(m, nn.Module) for m in trainable)
            for module in trainable:
                module.train()
        else:
            self.model.train()

    def set_eval(self):
        self.model.eval()

    def _to_device(self, data):
        if isinstance(data, (list, tuple)):
            return [self._to_device(d) for d in data]
        if isinstance(data, torch.Tensor):
            return data.to(self.device, non_blocking=True)
        if isinstance(data, dict):
            if 'input_ids_i32' in data:  # collate_fun(compact=True): int32 ids only, the rest is built on the GPU
                return expand_compact_inputs(data, self.device)
            return {k: self._to_device(v) for k, v in data.items()}
        if data is None:
            return data
        raise NotImplementedError(type(data))

    # ---- training -----------------------------------------------------------------------------
    def train(self, after_epoch_funcs=None):
        if self.train_dataloader is None:
            logger.warning('You have not specified train dataset, so you cannot run train method.')
            return
        after_epoch_funcs = [] if after_epoch_funcs is None else after_epoch_funcs
        for epoch_i in range(1, self.n_epochs + 1):
            self._train(epoch_i)
            for func in after_epoch_funcs:
                func(epoch_i)

    def train_micro_batch(self, inputs, labels, *, last_of_step):
        """trainer.py:276-280 for one micro-batch already on the device: forward, loss, backward
        (+ overlapped gradient all-reduce when this micro-batch completes an optimizer step).
        Returns the device tensor of the six loss scalars."""
        if self._loss_scale_cfg is None:
            self._loss_scale_cfg = self.loss.loss_config(loss_scale=1.0 / self.batch_split, device=self.device)
        lc = self._loss_scale_cfg
        if self.reducer is None or not last_of_step:
            _, losses = self.model.forward_loss_backward(inputs, labels, lc)
        else:
            _, losses = self.model.forward_loss_backward(inputs, labels, lc, last_phase=-1)
            eng = self.model.engine
            then = None
            if self.max_grad_norm and hasattr(self.optimizer, 'add_bucket_sumsq') and self.reducer.cuda:
                main = torch.cuda.current_stream(self.device)
                with torch.cuda.stream(self.reducer.stream):
                    self.reducer.stream.wait_stream(main)  # the previous step() has read the old sum
                    self.optimizer.begin_bucket_sumsq()
                then = self.optimizer.add_bucket_sumsq
            for first, last, bucket in self.reducer.phases():
                eng.backward(first, last)
                self.reducer.reduce_bucket(bucket, then=then)
            self.reducer.wait()
        return losses

    def optimizer_step(self):
        """trainer.py:285-290."""
        self._clip_grad_norm()
        self.optimizer.step()
        self.optimizer.zero_grad()
        if self.scheduler is not None:
            self.scheduler.step()

    @time_profiler
    def _train(self, epoch_i):
        from tqdm.auto import tqdm
        self.set_train()
        self.optimizer.zero_grad()
        avg_meters = defaultdict(AverageMeter)
        readback = _LossReadback(self.device)
        tqdm_data = tqdm(self.train_dataloader, desc=f'Train (epoch #{epoch_i} / {self.n_epochs})',
                         disable=self.local_rank not in (-1, 0))
        for i, (inputs, labels) in enumerate(tqdm_data):
            inputs, labels = self._to_device((inputs, labels))
            last = (i + 1) % self.batch_split == 0
            readback.push(self.train_micro_batch(inputs, labels, last_of_step=last))
            vals = readback.pop_previous()  # one iteration late: never waits for the GPU
            if vals is not None:
                if vals[5] != vals[5]:
                    # NaN: either divergence (the reference trains on) or token ids outside the embedding tables,
                    # which the kernels turn into a NaN loss where torch raises IndexError — ask and raise
                    self.model.engine.check_inputs()
                avg_meters.update(zip(_LOSS_KEYS, vals))
            avg_meters['lr'] = self._get_lr()
            if last:
                self.optimizer_step()
                self._update_writer(avg_meters, prefix='train')
                self.global_step += 1
                if self.debug:
                    logger.info('Training was interrupted because of debug mode.')
                    break
            tqdm_data.set_postfix_str(self._get_console_str(avg_meters))
        vals = readback.pop_last()
        if vals is not None:
            avg_meters.update(zip(_LOSS_KEYS, vals))
        self.last_train_meters = dict(avg_meters)

    # ---- evaluation ---------------------------------------------------------------------------
    def test(self, epoch_i, *, callbacks=None):
        if self.local_rank in [0, -1]:
            if self.test_dataloader is None:
                logger.warning('You have not specified test dataset, so you cannot run test method.')
                return
            if callbacks is not None and not isinstance(callbacks, (list, tuple)):
                callbacks = tuple(callbacks)
            assert all(isinstance(c, TestCallback) for c in (callbacks or ()))
            with torch.no_grad():
                self._test(epoch_i, callbacks=callbacks)
        if self.local_rank != -1 and torch.distributed.is_initialized():
            logger.warning('Waiting till validation ends in main process..')
            torch.distributed.barrier()

    @torch.no_grad()
    @time_profiler
    def _test(self, epoch_i, *, callbacks=None):
        from tqdm.auto import tqdm
        self.set_eval()
        avg_meters = defaultdict(AverageMeter)
        tqdm_data = tqdm(self.test_dataloader, desc=f'Test (epoch #{epoch_i} / {self.n_epochs})')
        for i, (inputs, labels) in enumerate(tqdm_data):
            inputs, labels = self._to_device((inputs, labels))
            preds = self.model(**inputs)
            self.loss(preds, labels, avg_meters=avg_meters)
            for callback in callbacks or ():
                callback.at_iteration_end(preds, labels, avg_meters)
            tqdm_data.set_postfix_str(self._get_console_str(avg_meters))
            if self.debug and i >= 10:
                logger.info('Test was interrupted because of debug mode.')
                break
        for callback in callbacks or ():
            callback.at_epoch_end(avg_meters, self)
        self._update_writer(avg_meters, prefix='test')
        metrics = {k: v() if isinstance(v, AverageMeter) else v for k, v in avg_meters.items()}
        logger.info(f'Test metrics after epoch {epoch_i} - {self._get_console_str(metrics)}')
        self.last_test_metrics = metrics

    # ---- checkpoints (wire format of trainer.py:355-403) --------------------------------------
    def save_state_dict(self, path_):
        if self.local_rank not in [-1, 0]:
            return
        if self.debug:
            logger.info(f'Model was not saved to {path_} because of debug mode.')
            return
        state = {'model': {k: v.detach().clone() for k, v in self.model.state_dict().items()},
                 'optimizer': self.optimizer.state_dict(),
                 'scheduler': self.scheduler.state_dict() if self.scheduler is not None else None,
                 'global_step': self.global_step}
        os.makedirs(os.path.dirname(str(path_)) or '.', exist_ok=True)
        torch.save(state, path_)
        logger.info(f'State dict was saved to {path_}.')

    def load_state_dict(self, path_):
        if not os.path.exists(path_):
            logger.warning(f'Checkpoint {path_} does not exist, so checkpoint was not loaded.')
            return
        state = torch.load(path_, map_location=self.device, weights_only=False)  # the recipe's own, trusted pickles
        # checkpoints written by the reference carry transformers' embedding buffers (position_ids): not parameters
        self.model.load_state_dict({k: v for k, v in state['model'].items()
                                    if not k.endswith(('embeddings.position_ids', 'embeddings.token_type_ids'))})
        self.global_step = state['global_step']
        logger.info(f'Model weights were loaded from {path_} checkpoint.')
        if not self.drop_optimizer:
            self.optimizer.load_state_dict(state['optimizer'])
            if self.scheduler is not None and state.get('scheduler') is not None:
                self.scheduler.load_state_dict(state['scheduler'])
            logger.info(f'Optimizer and scheduler also were restored from {path_} checkpoint.')
