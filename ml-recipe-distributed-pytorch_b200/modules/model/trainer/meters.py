"""Running means and per-class average precision (reference meters.py:10-58); host-side."""
from collections import defaultdict

import numpy as np


class AverageMeter:
    def __init__(self):
        self._n, self._avg = 0, 0

    def __call__(self):
        return self._avg

    def update(self, value):
        self._n += 1
        self._avg += (value - self._avg) / self._n


class APMeter:
    def __init__(self):
        self.reset()

    def __call__(self):
        from sklearn import metrics
        return metrics.average_precision_score(self.true_labels, self.pred_probas)

    def update(self, pred_probas, true_labels):
        self.pred_probas.extend(pred_probas)
        self.true_labels.extend(true_labels)

    def reset(self):
        self.pred_probas, self.true_labels = [], []


class MAPMeter:
    def __init__(self):
        self.reset()

    def __call__(self):
        out = {k: v() for k, v in self.aps_dict.items()}
        out['map'] = np.mean(list(out.values()))
        return out

    def update(self, keys, pred_probas, true_labels):
        assert len(keys) == pred_probas.shape[-1]
        for i, key in enumerate(keys):
            self.aps_dict[key].update(pred_probas[:, i], true_labels == i)

    def reset(self):
        self.aps_dict = defaultdict(APMeter)
