"""Synthetic full-length QA items: [CLS] question(64) [SEP] document [SEP], fixed labels.
Same distribution and labels as the reference's DummyDataset (dummy_dataset.py:6-51); it is the
input generator of every BASELINE configuration."""
import numpy as np

from .items import DatasetItem


class DummyDataset:
    def __init__(self, tokenizer, *args, max_seq_len=384, max_question_len=64, dataset_len=10000, **kwargs):
        self.tokenizer = tokenizer
        self.dataset_len = dataset_len
        self.max_seq_len = max_seq_len
        self.max_question_len = max_question_len
        self.w_ids = None if tokenizer is None else [tokenizer.pad_token_id, tokenizer.sep_token_id,
                                                     tokenizer.cls_token_id]

    def __len__(self):
        return self.dataset_len

    def _random_ids(self, n):
        assert self.w_ids is not None, f'Dataset {type(self).__name__} was initialized with None tokenizer.'
        ids = np.random.randint(1, len(self.tokenizer), n)
        ids[np.isin(ids, self.w_ids)] = self.tokenizer.unk_token_id
        return ids.tolist()

    def __getitem__(self, *args):
        question = self._random_ids(self.max_question_len)
        document = self._random_ids(self.max_seq_len - self.max_question_len - 3)
        tk = self.tokenizer
        return DatasetItem(input_ids=[tk.cls_token_id] + question + [tk.sep_token_id] + document + [tk.sep_token_id],
                           start_id=0, end_id=self.max_seq_len - 1, label_id=0, example_id='None',
                           start_position=0, end_position=1)
