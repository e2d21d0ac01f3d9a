"""Inference entry point (reference modules/validate.py:29-63): restore a checkpoint, run the
Predictor over a chunk dataset.  The reference's ChunkDataset reads the preprocessed Natural
Questions dump, which is not shipped; `main` therefore takes the dataset as an argument and the
command line builds the synthetic chunk dataset of BASELINE config 5.

The reference validates in one process.  Under torchrun (WORLD_SIZE > 1) this entry point shards the
DOCUMENTS across the ranks (one process per GPU): all chunks of a document are scored on one rank, so
the per-document best-candidate rule stays local, and the winners are gathered on rank 0."""
import os
import sys

import numpy as np
import torch

if __package__ in (None, ''):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from __graft_entry__ import load_package
    load_package()
    __package__ = 'bqa_b200.modules'

from .init import init_collate_fun, init_model  # noqa: E402
from .model.dataset import ChunkItem  # noqa: E402
from .model.inference.predictor import DocumentShard, Predictor, gather_predictions  # noqa: E402
from .model.utils.parser import get_model_parser, get_params, get_predictor_parser  # noqa: E402
from .utils import get_logger, show_params  # noqa: E402


class SyntheticChunkDataset:
    """Documents of random length cut into windows of max_seq_len with stride doc_stride; one
    dataset item = all chunks of one document (the shape of ChunkDataset items,
    validation_dataset.py:42-319).  Used where the real dump is unavailable."""

    def __init__(self, n_docs, *, max_seq_len=512, max_question_len=64, doc_stride=128, vocab=30522, seed=1234,
                 max_doc_len=2048):
        rng = np.random.RandomState(seed)
        self.docs = []
        for d in range(n_docs):
            qlen = int(rng.randint(5, max_question_len + 1))
            dlen = int(rng.randint(64, max_doc_len))
            q = rng.randint(1000, vocab, qlen).tolist()
            doc = rng.randint(1000, vocab, dlen).tolist()
            window = max_seq_len - qlen - 3
            chunks, start = [], 0
            while True:
                piece = doc[start:start + window]
                ids = [101] + q + [102] + piece + [102]
                chunks.append(ChunkItem(item_id=f'doc{d}', input_ids=ids, question_len=qlen, chunk_start=start,
                                        chunk_end=start + len(piece)))
                if start + window >= dlen:
                    break
                start += doc_stride
            self.docs.append(chunks)

    def __len__(self):
        return len(self.docs)

    def __getitem__(self, i):
        return self.docs[i]


def main(params, model_params, dataset=None):
    show_params(model_params, 'model')
    show_params(params, 'predictor')
    if not (torch.cuda.is_available() and params.gpu):
        raise RuntimeError('validate needs --gpu with a CUDA device (sm_100a kernels only)')
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    if world > 1:
        torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', rank)))
        if not torch.distributed.is_initialized():
            torch.distributed.init_process_group(backend='nccl')
    device = torch.device('cuda', torch.cuda.current_device())
    checkpoint = params.checkpoint if params.checkpoint and os.path.exists(params.checkpoint) else None
    model, tokenizer = init_model(model_params, checkpoint=checkpoint, device=device)
    if dataset is None:
        dataset = SyntheticChunkDataset(64, max_seq_len=params.max_seq_len, max_question_len=params.max_question_len,
                                        doc_stride=params.doc_stride)
    if world > 1:
        dataset = DocumentShard(dataset, rank, world)
    predictor = Predictor(model, device, collate_fun=init_collate_fun(tokenizer, return_items=True),
                          batch_size=params.batch_size, n_jobs=params.n_jobs, buffer_size=params.buffer_size,
                          limit=params.limit)
    predictor(dataset)
    gather_predictions(predictor)
    return predictor


if __name__ == '__main__':
    _, (params, model_params) = get_params((get_predictor_parser, get_model_parser))
    logger = get_logger(logger_name='validate')
    main(params, model_params)
