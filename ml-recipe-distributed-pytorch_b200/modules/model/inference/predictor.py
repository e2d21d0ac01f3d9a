"""Batched inference + span decoding with the reference Predictor's interface
(modules/model/inference/predictor.py:22-144).

The reference copies all five outputs ([B,S] x2, [B,5], [B] x2) to the host after every batch
and runs torch.max / softmax there (predictor.py:111-120); here bqa_decode produces the arg max
indices, class ids and scores on the device and only those ~7 small vectors are copied.  The
per-document candidate selection (predictor.py:63-87) is unchanged host logic.
"""
import dataclasses
import logging
from collections import defaultdict
from dataclasses import dataclass

import torch

from ..trainer.callback import device_decode

logger = logging.getLogger(__name__)


@dataclass
class PredictorCandidate:
    start_id: int
    end_id: int
    start_reg: float
    end_reg: float
    label: int


class DocumentShard:
    """Every `world`-th document of a chunk dataset, starting at `rank`.  A dataset item is ALL chunks of one
    document (validation_dataset.py:42-319), so every chunk of a document is scored on the same rank and the
    per-document best-candidate rule of the Predictor (predictor.py:63-87) needs no communication; only the
    per-document winners are gathered at the end (`gather_predictions`)."""

    def __init__(self, dataset, rank, world):
        self.dataset = dataset
        self.indices = list(range(rank, len(dataset), world))

    def __len__(self):
        return len(self.indices)

    def __getitem__(self, i):
        return self.dataset[self.indices[i]]


def gather_predictions(predictor, dst=0, group=None):
    """Data-parallel validation: merges the per-document winners of every rank's Predictor into rank `dst`'s
    `scores` / `candidates` / `items` (documents are disjoint across ranks under DocumentShard; should an item_id
    appear twice, the higher score wins, which is the Predictor's own rule).  Returns the number of documents on
    `dst`, 0 elsewhere.  No-op without an initialised process group."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return len(predictor.candidates)
    payload = {k: (predictor.scores[k], dataclasses.astuple(predictor.candidates[k]), predictor.items.get(k))
               for k in predictor.candidates}
    gathered = [None] * dist.get_world_size(group)
    dist.all_gather_object(gathered, payload, group=group)
    if dist.get_rank(group) != dst:
        return 0
    for part in gathered:
        for k, (score, cand, item) in part.items():
            if k in predictor.candidates and predictor.scores[k] >= score:
                continue
            predictor.scores[k] = score
            predictor.candidates[k] = PredictorCandidate(*cand)
            predictor.items[k] = item
    return len(predictor.candidates)


class Predictor:
    def __init__(self, model, device, *, batch_size=256, n_jobs=16, collate_fun=None, buffer_size=4096, limit=None):
        self.model = model
        self.device = torch.device(device)
        self.model.to(self.device)
        self.scores = defaultdict(int)
        self.candidates = {}
        self.items = {}
        self.batch_size = batch_size
        self.n_jobs = n_jobs
        self.collate_fun = collate_fun
        self.buffer_size = buffer_size
        self.limit = limit
        self.dump = None
        logger.info(f'Predictor uses {self.device} device. Batch size: {self.batch_size}. Set limit: {self.limit}.')

    def _is_valid(self, item, score, start_id, end_id):
        assert score >= 0
        if start_id > end_id:
            return False
        if start_id < item.question_len + 2:
            return False
        if self.scores[item.item_id] > score:
            return False
        return True

    def _update_candidates(self, scores, start_ids, end_ids, start_regs, end_regs, labels, items):
        for score, s, e, sr, er, label, item in zip(scores, start_ids, end_ids, start_regs, end_regs, labels, items):
            if self._is_valid(item, score, s, e):
                self.scores[item.item_id] = score
                self.candidates[item.item_id] = PredictorCandidate(start_id=s, end_id=e, start_reg=sr, end_reg=er,
                                                                   label=label)
                self.items[item.item_id] = item

    def predict_batch(self, inputs):
        """Forward + decode of one collated batch; returns host numpy vectors."""
        inputs = {k: v.to(self.device, non_blocking=True) for k, v in inputs.items()}
        preds = self.model(**inputs)
        dec = device_decode(preds)
        ints = torch.stack([dec['start_ids'], dec['end_ids'], dec['cls_ids']]).cpu().numpy()
        flts = torch.stack([dec['scores'], preds['start_reg'].float(), preds['end_reg'].float()]).cpu().numpy()
        # the copies above synchronised the stream: reading the input-error flag costs nothing more.  The
        # reference's nn.Embedding raises IndexError for ids outside the tables; so does this.
        self.model.engine.check_inputs()
        return {'scores': flts[0], 'start_ids': ints[0], 'end_ids': ints[1], 'cls_ids': ints[2],
                'start_reg': flts[1], 'end_reg': flts[2]}

    def _batches(self, dataset):
        """Dataset items may be single chunks or lists of chunks (ChunkDataset returns every chunk of
        a document); flattened into batches of `batch_size` like the reference's ListDataloader."""
        buf = []
        for idx in range(len(dataset)):
            item = dataset[idx]
            buf.extend(item if isinstance(item, (list, tuple)) else [item])
            while len(buf) >= self.batch_size:
                yield self.collate_fun(buf[:self.batch_size])
                buf = buf[self.batch_size:]
        if buf:
            yield self.collate_fun(buf)

    @torch.no_grad()
    def __call__(self, dataset, *, save_dump=False):
        self.model.eval()
        if save_dump:
            self.dump = []
        for batch_i, (inputs, labels, items) in enumerate(self._batches(dataset)):
            out = self.predict_batch(inputs)
            self._update_candidates(out['scores'], out['start_ids'], out['end_ids'], out['start_reg'],
                                    out['end_reg'], out['cls_ids'], items)
            if save_dump:
                self.dump.append((out['scores'], out['start_ids'], out['end_ids'], out['cls_ids'], items))
            if self.limit is not None and batch_i >= self.limit:
                break
