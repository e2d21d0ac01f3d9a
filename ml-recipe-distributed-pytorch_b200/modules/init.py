"""Factories with the reference's names and arguments (modules/init.py:18-205): loss set,
model + tokenizer stand-in, parameter groups + optimizer, datasets, collate function."""
import functools
import logging
import os
from collections import defaultdict

import torch
import torch.nn as nn

from .model.dataset import DummyDataset, RawPreprocessor, collate_fun
from .model.model import (BertForQuestionAnswering, FocalLossWithLogits, LabelSmoothingLossWithLogits,
                          WeightedLoss)
from .model.trainer.optim import AdaMod, FusedAdamW

logger = logging.getLogger(__name__)


class DummyTokenizer:
    """What DummyDataset and collate_fun need when no vocab file is shipped (the reference falls
    back to downloading one, init.py:65-73, which is impossible offline): the special ids of the
    model family (BERT: [PAD] 0, [UNK] 100, [CLS] 101, [SEP] 102; RoBERTa: <s> 0, <pad> 1, </s> 2, <unk> 3)."""
    pad_token_id, unk_token_id, cls_token_id, sep_token_id = 0, 100, 101, 102

    def __init__(self, model_name='bert', vocab_size=30522):
        self.model_name = model_name
        self._n = vocab_size
        if model_name == 'roberta':
            self.pad_token_id, self.unk_token_id, self.cls_token_id, self.sep_token_id = 1, 3, 0, 2

    def __len__(self):
        return self._n


def init_loss(params, train_weights):
    def _wght(name_):
        return getattr(params, name_, 1)

    if params.loss == 'ce':
        class_loss = nn.CrossEntropyLoss(weight=train_weights['label_weights'])
    elif params.loss == 'focal':
        class_loss = FocalLossWithLogits(alpha=params.focal_alpha, gamma=params.focal_gamma)
    elif params.loss == 'smooth':
        class_loss = LabelSmoothingLossWithLogits(n_classes=len(RawPreprocessor.labels2id),
                                                  smoothing=params.smooth_alpha)
    else:
        raise NotImplementedError
    logger.info(f'Used loss function for classification: {type(class_loss).__name__}.')
    return WeightedLoss({'start_class': (nn.CrossEntropyLoss(ignore_index=-1), _wght('w_start')),
                         'end_class': (nn.CrossEntropyLoss(ignore_index=-1), _wght('w_end')),
                         'start_reg': (nn.MSELoss(), _wght('w_start_reg')),
                         'end_reg': (nn.MSELoss(), _wght('w_end_reg')),
                         'cls': (class_loss, _wght('w_cls'))})


# buffers that transformers registers on the embeddings (persistent in the reference's pinned 3.5.1): present in
# checkpoints written by the reference, not parameters of the drop-in
_EMBEDDING_BUFFERS = ('embeddings.position_ids', 'embeddings.token_type_ids')


def strip_reference_buffers(model_state):
    return {k: v for k, v in model_state.items() if not k.endswith(_EMBEDDING_BUFFERS)}


def _load_checkpoint(model, checkpoint, *, device=torch.device('cpu')):
    if checkpoint is not None:
        # the recipe's own checkpoints (trainer.py:355-379): trusted pickles, as with the reference's torch 1.3
        state_dict = torch.load(checkpoint, map_location=device, weights_only=False)
        model.load_state_dict(state_dict['model'], strict=False)
        logger.info(f'Model checkpoint was restored from {checkpoint}.')


def init_model(model_params, *, checkpoint=None, device=torch.device('cpu'), bpe_dropout=None):
    model_params.model_name = model_params.model.split('-')[0]
    if model_params.model_name not in ('bert', 'roberta'):
        raise NotImplementedError(f'{model_params.model}: the reference knows bert and roberta encoders (model.py:9-10)')
    vocab_file = getattr(model_params, 'vocab_file', None)
    if vocab_file is not None and os.path.exists(vocab_file):
        with open(vocab_file) as f:
            vocab_size = sum(1 for _ in f)
    else:
        logger.warning('No vocab file: using the special-token stand-in tokenizer (dummy data only).')
        vocab_size = 50265 if model_params.model_name == 'roberta' else 30522
    tokenizer = DummyTokenizer(model_params.model_name, vocab_size)
    model = BertForQuestionAnswering(model_params, num_labels=len(RawPreprocessor.labels2id))
    model.to(device)
    if checkpoint is not None:
        _load_checkpoint(model, checkpoint, device=device)
    return model, tokenizer


def _get_optimized_parameters(params, model):
    if params.finetune:
        model.eval()  # reference init.py:92: freezes dropout while fine-tuning heads
        modules, named = [], []
        for flag, mods in (('finetune_transformer', [model.transformer]),
                           ('finetune_position', [model.position_outputs]),
                           ('finetune_position_reg', [model.reg_start, model.reg_end]),
                           ('finetune_class', [model.classifier])):
            if getattr(params, flag):
                for m in mods:
                    modules.append(m)
                    named.extend(m.named_parameters())  # names relative to the sub-module, as in init.py:99-113
        if not modules:
            raise AttributeError('Specify at least one module for fine-tuning.')
    else:
        modules, named = None, list(model.named_parameters())
    no_decay = ['bias', 'LayerNorm.bias', 'LayerNorm.weight']
    groups = [{'params': [p for n, p in named if not any(nd in n for nd in no_decay)],
               'weight_decay': params.weight_decay},
              {'params': [p for n, p in named if any(nd in n for nd in no_decay)], 'weight_decay': 0.0}]
    return modules, groups


def init_optimizer(params, model):
    modules, groups = _get_optimized_parameters(params, model)
    if params.optimizer == 'adam':
        optimizer = FusedAdamW(groups, model, lr=params.lr, correct_bias=False)
    else:
        optimizer = AdaMod(groups, model, lr=params.lr)
    logger.info(f'Used optimizer: {type(optimizer).__name__}.')
    if modules is not None:
        model.list_of_trainable_modules = modules
    return optimizer


def init_datasets(params, *, tokenizer=None, clear=False):
    weights = defaultdict(lambda: None)
    if not params.dummy_dataset:
        raise NotImplementedError('real-data pipeline (reference init.py:161-177) is outside the hot path; '
                                  'set dummy_dataset=True')
    logger.warning('Dummy dataset is used to train model.')
    kw = dict(tokenizer=tokenizer, max_seq_len=params.max_seq_len, max_question_len=params.max_question_len)
    train_dataset = DummyDataset(**kw)
    test_dataset = DummyDataset(**kw) if getattr(params, 'local_rank', -1) in [-1, 0] else None
    return train_dataset, test_dataset, weights


def init_collate_fun(tokenizer, return_items=False, compact=False):
    """compact=True: batches carry int32 ids only and Trainer._to_device builds the model inputs on the GPU."""
    return functools.partial(collate_fun, tokenizer=tokenizer, return_items=return_items, compact=compact)
