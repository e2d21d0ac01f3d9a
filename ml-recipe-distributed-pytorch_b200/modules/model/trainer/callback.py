"""Test-time callbacks (reference callback.py:12-108).  Arg max runs in the decode kernel on the
device; only index vectors (and the [B, 5] class logits for AP) cross to the host."""
import ctypes
import logging
import math

import torch

from .... import _lib as L
from .meters import AverageMeter, MAPMeter

logger = logging.getLogger(__name__)


def device_decode(preds):
    """bqa_decode over the output dict -> dict of device tensors (predictor.py:114-120 semantics)."""
    start = preds['start_class'].detach().float().contiguous()
    end = preds['end_class'].detach().float().contiguous()
    cls = preds['cls'].detach().float().contiguous()
    B, S = start.shape
    C = cls.shape[1]
    dev = start.device
    idx = torch.empty(3, B, dtype=torch.int64, device=dev)
    val = torch.empty(4, B, dtype=torch.float32, device=dev)
    L.check(L.lib.bqa_decode(start.data_ptr(), end.data_ptr(), cls.data_ptr(), idx[0].data_ptr(), idx[1].data_ptr(),
                             idx[2].data_ptr(), val[0].data_ptr(), val[1].data_ptr(), val[2].data_ptr(),
                             val[3].data_ptr(), B, S, C, L.current_stream_ptr()), "bqa_decode")
    return {'start_ids': idx[0], 'end_ids': idx[1], 'cls_ids': idx[2], 'start_logits': val[0], 'end_logits': val[1],
            'cls_probas': val[2], 'scores': val[3]}


class TestCallback:
    def at_iteration_end(self, preds, labels, avg_meters):
        self._at_iteration_end(preds, labels, avg_meters)

    def _at_iteration_end(self, *args):
        raise NotImplementedError

    def at_epoch_end(self, avg_meters, trainer):
        self._at_epoch_end(avg_meters, trainer)
        self._reset()

    def _at_epoch_end(self, *args):
        raise NotImplementedError

    def _reset(self):
        pass


class AccuracyCallback(TestCallback):
    keys = ['start_class', 'end_class', 'cls']

    def _at_iteration_end(self, preds, labels, avg_meters):
        dec = device_decode(preds)
        pred = torch.stack([dec['start_ids'], dec['end_ids'], dec['cls_ids']]).cpu()
        true = [labels[k].detach().cpu() for k in self.keys]
        for name, p, t, ignore in (('s_acc', pred[0], true[0], -1), ('e_acc', pred[1], true[1], -1),
                                   ('c_acc', pred[2], true[2], None)):
            keep = torch.ones_like(t, dtype=torch.bool) if ignore is None else t != ignore
            if bool(keep.any()):
                avg_meters[name].update(float((p[keep] == t[keep]).float().mean()))

    def _at_epoch_end(self, *args):
        pass


class MAPCallback(TestCallback):
    key = 'cls'

    def __init__(self, metric_keys):
        self._metric_keys = metric_keys
        self._reset()

    def _at_iteration_end(self, preds, labels, *args):
        logits = preds[self.key].detach().float().cpu()
        self.map_meter.update(keys=self._metric_keys, pred_probas=torch.softmax(logits, dim=-1).numpy(),
                              true_labels=labels[self.key].detach().cpu().numpy())

    def _at_epoch_end(self, avg_meters, *args):
        avg_meters.update(self.map_meter())

    def _reset(self):
        self.map_meter = MAPMeter()


class SaveBestCallback(TestCallback):
    def __init__(self, params):
        self.params = params
        self.metric = params.best_metric
        self.best_order = params.best_order
        self.value = 1e10 * (-1 if self.best_order == '>' else 1)

    def _at_iteration_end(self, *args):
        pass

    def _at_epoch_end(self, avg_meters, trainer):
        metrics = {k: v() if isinstance(v, AverageMeter) else v for k, v in avg_meters.items()}
        value = metrics.get(self.metric)
        if value is None or math.isnan(value):
            logger.warning(f'Trainer metrics do not contain metric {self.metric}.')
            return
        better = value > self.value if self.best_order == '>' else value < self.value
        if better:
            self.value = value
            trainer.save_state_dict(self.params.dump_dir / self.params.experiment_name / 'best.ch')
            logger.info(f'Best value of {self.metric} was achieved after training step {trainer.global_step} '
                        f'and equals to {self.value:.3f}')
        else:
            logger.info(f'Best value {self.value:.3f} of {self.metric} was not beaten with {value:.3f}')
