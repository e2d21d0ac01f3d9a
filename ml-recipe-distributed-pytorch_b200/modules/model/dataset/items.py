"""Input ABI of the step: item records and the collate function that turns them into the
tensors `BertForQuestionAnswering.forward` and `WeightedLoss` consume.

Mirrors the reference's DatasetItem / collate_fun (modules/model/dataset/split_dataset.py:191-199,
480-520) and the field contract of ChunkItem (validation_dataset.py:15-39).  The reference's
real-data pipeline (RawPreprocessor / SplitDataset / ChunkDataset: json lines, nltk sentence
splitting, tokenizers) is outside the hot path and its dataset is not shipped; only the label
table of RawPreprocessor is needed here (split_dataset.py:51-52).
"""
from dataclasses import dataclass, field
from typing import List

import numpy as np
import torch

LABELS2ID = {k: i for i, k in enumerate(['yes', 'no', 'short', 'long', 'unknown'])}
ID2LABELS = {i: k for k, i in LABELS2ID.items()}


class RawPreprocessor:
    """Label tables only; preprocessing of the Natural Questions dump is out of scope."""
    labels2id = LABELS2ID
    id2labels = ID2LABELS

    def __init__(self, *a, **k):
        raise NotImplementedError("real-data preprocessing (reference split_dataset.py:50-188) is outside the "
                                  "B200 hot path; run with dummy_dataset=True or feed items directly")


@dataclass
class DatasetItem:
    example_id: str
    input_ids: List[int]
    start_id: int
    end_id: int
    label_id: int
    start_position: float
    end_position: float


@dataclass
class ChunkItem:
    """What Predictor needs from a validation chunk: item_id, question_len, input_ids (+ labels)."""
    item_id: str
    input_ids: List[int]
    question_len: int
    start_id: int = -1
    end_id: int = -1
    label_id: int = 0
    start_position: float = 0.0
    end_position: float = 0.0
    true_text: str = ''
    true_question: str = ''
    true_label: int = 0
    true_start: int = -1
    true_end: int = -1
    chunk_start: int = 0
    chunk_end: int = 0
    t2o: List[int] = field(default_factory=list)


def collate_fun(items, tokenizer, return_items=False, pin_memory=False, compact=False):
    """Pads to the batch maximum; token_type 0 up to and including the first [SEP], 1 after
    (BERT only); attention_mask = ids > 0; labels as int64 / fp32 vectors.

    compact=True ships only int32 token ids (`input_ids_i32`, plus the two scalars the device needs):
    `expand_compact_inputs` / `Trainer._to_device` rebuild input_ids, token_type_ids and
    attention_mask on the GPU (bqa_prepare_inputs), a quarter of the host-to-device bytes and no
    per-row Python loop over positions."""
    n = len(items)
    max_len = max(len(it.input_ids) for it in items)
    if compact:
        return _collate_compact(items, tokenizer, n, max_len, return_items, pin_memory)
    tokens = np.full((n, max_len), tokenizer.pad_token_id, dtype=np.int64)
    bert = tokenizer.model_name == 'bert'
    token_type = np.full((n, max_len), 1 if bert else 0, dtype=np.int64)
    sep = tokenizer.sep_token_id
    for i, it in enumerate(items):
        row = np.asarray(it.input_ids, dtype=np.int64)
        tokens[i, :len(row)] = row
        if bert:
            first_sep = int(np.flatnonzero(row == sep)[0])  # ValueError-equivalent if absent, like list.index
            token_type[i, :first_sep + 1] = 0
    mask = tokens > 0
    inputs = {'input_ids': torch.from_numpy(tokens), 'attention_mask': torch.from_numpy(mask),
              'token_type_ids': torch.from_numpy(token_type)}
    labels = {'start_class': torch.tensor([it.start_id for it in items], dtype=torch.int64),
              'end_class': torch.tensor([it.end_id for it in items], dtype=torch.int64),
              'start_reg': torch.tensor([it.start_position for it in items], dtype=torch.float32),
              'end_reg': torch.tensor([it.end_position for it in items], dtype=torch.float32),
              'cls': torch.tensor([it.label_id for it in items], dtype=torch.int64)}
    if pin_memory and torch.cuda.is_available():
        inputs = {k: v.pin_memory() for k, v in inputs.items()}
        labels = {k: v.pin_memory() for k, v in labels.items()}
    if return_items:
        return [inputs, labels, items]
    return [inputs, labels]


def _labels(items):
    return {'start_class': torch.tensor([it.start_id for it in items], dtype=torch.int64),
            'end_class': torch.tensor([it.end_id for it in items], dtype=torch.int64),
            'start_reg': torch.tensor([it.start_position for it in items], dtype=torch.float32),
            'end_reg': torch.tensor([it.end_position for it in items], dtype=torch.float32),
            'cls': torch.tensor([it.label_id for it in items], dtype=torch.int64)}


def _collate_compact(items, tokenizer, n, max_len, return_items, pin_memory):
    bert = tokenizer.model_name == 'bert'
    sep = tokenizer.sep_token_id
    tokens = np.full((n, max_len), tokenizer.pad_token_id, dtype=np.int32)
    for i, it in enumerate(items):
        row = np.asarray(it.input_ids, dtype=np.int32)
        if bert and not (row == sep).any():
            raise ValueError(f'{sep} is not in list')  # the reference's row.index(sep_token_id)
        tokens[i, :len(row)] = row
    # no_padding: every position holds a real token (ids > 0), so attention_mask would be all True; the model
    # then gets attention_mask=None (same result) and the attention kernels take their mask-free variants
    inputs = {'input_ids_i32': torch.from_numpy(tokens), 'sep_token_id': int(sep), 'type_coef': 1 if bert else 0,
              'no_padding': bool((tokens > 0).all())}
    labels = _labels(items)
    if pin_memory and torch.cuda.is_available():
        inputs['input_ids_i32'] = inputs['input_ids_i32'].pin_memory()
        labels = {k: v.pin_memory() for k, v in labels.items()}
    if return_items:
        return [inputs, labels, items]
    return [inputs, labels]


def expand_compact_inputs(inputs, device):
    """{'input_ids_i32', 'sep_token_id', 'type_coef'} -> the model's input dict on `device`, built by
    bqa_prepare_inputs (collate_fun's token-type / mask rules, include/bqa.h)."""
    import ctypes

    from .... import _lib as L
    ids32 = inputs['input_ids_i32'].to(device, non_blocking=True)
    if ids32.device.type != 'cuda':
        raise RuntimeError('compact inputs are expanded by a CUDA kernel: there is no CPU path')
    B, S = ids32.shape
    ids = torch.empty(B, S, dtype=torch.int64, device=device)
    tt = torch.empty(B, S, dtype=torch.int64, device=device)
    mask = torch.empty(B, S, dtype=torch.bool, device=device)
    with torch.cuda.device(ids32.device):
        L.check(L.lib.bqa_prepare_inputs(ids32.data_ptr(), B, S, int(inputs['sep_token_id']), int(inputs['type_coef']),
                                         ids.data_ptr(), tt.data_ptr(), mask.data_ptr(), None,
                                         L.current_stream_ptr()), 'bqa_prepare_inputs')
    return {'input_ids': ids, 'attention_mask': None if inputs.get('no_padding', False) else mask,
            'token_type_ids': tt}
