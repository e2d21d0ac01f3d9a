"""Training entry point with the reference's command line (modules/train.py:125-167):

    python -m torch.distributed.run --nproc-per-node N train.py -c config/test_bert.cfg ...   (or)
    python train.py -c config/test_bert.cfg --local_rank 0 ...                               (1 GPU)

One process per GPU.  The reference spawns its workers itself with mp.spawn and a tcp://
rendezvous (train.py:133-148); here the launcher is torchrun (RANK/LOCAL_RANK/WORLD_SIZE from
the environment) and, for compatibility, the reference's explicit --local_rank / --dist_* flags
are honoured when the environment variables are absent.
"""
import functools
import logging
import os
import sys
from datetime import datetime

import torch

if __package__ in (None, ''):  # executed as a script: make the package importable
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from __graft_entry__ import load_package
    load_package()
    import importlib
    _pkg = importlib.import_module('bqa_b200.modules')
    __package__ = 'bqa_b200.modules'

from .init import init_collate_fun, init_datasets, init_loss, init_model, init_optimizer  # noqa: E402
from .model.dataset import RawPreprocessor  # noqa: E402
from .model.trainer.callback import AccuracyCallback, MAPCallback, SaveBestCallback  # noqa: E402
from .model.trainer.trainer import Trainer  # noqa: E402
from .model.utils.parser import get_model_parser, get_params, get_trainer_parser, write_config_file  # noqa: E402
from .utils import get_logger, set_seed, show_params  # noqa: E402


def run_worker(device, params, model_params):
    logger = get_logger(level=logging.INFO if params.local_rank in [-1, 0] else logging.WARN,
                        filename=params.log_file if params.local_rank in [-1, 0] else None, filemode='a',
                        logger_name='train', debug=params.debug)
    logger.warning(f'Process with local_rank: {params.local_rank}. Used device: {device}.')
    model, tokenizer = init_model(model_params, device=device, bpe_dropout=params.bpe_dropout)
    optimizer = init_optimizer(params, model)
    train_dataset, test_dataset, train_weights = init_datasets(params, tokenizer=tokenizer, clear=False)
    if params.local_rank != -1 and torch.distributed.is_initialized():
        torch.distributed.barrier()
    loss = init_loss(params, train_weights)
    trainer = Trainer(model=model, loss=loss, collate_fun=init_collate_fun(tokenizer), optimizer=optimizer,
                      train_dataset=train_dataset, test_dataset=test_dataset,
                      writer_dir=params.dump_dir / f'board/{params.experiment_name}', device=device,
                      local_rank=params.local_rank, gpu_id=None, sync_bn=params.sync_bn, n_epochs=params.n_epochs,
                      train_batch_size=params.train_batch_size, test_batch_size=params.test_batch_size,
                      batch_split=params.batch_split, n_jobs=params.n_jobs, warmup_coef=params.warmup_coef,
                      max_grad_norm=params.max_grad_norm, apex_level=params.apex_level,
                      apex_verbosity=params.apex_verbosity, apex_loss_scale=params.apex_loss_scale,
                      train_weights=train_weights, drop_optimizer=params.drop_optimizer, debug=params.debug)
    if params.last is not None:
        trainer.load_state_dict(params.last)

    def save_last(*args, **kwargs):
        trainer.save_state_dict(params.dump_dir / params.experiment_name / 'last.ch')

    def save_each(epoch_i):
        trainer.save_state_dict(params.dump_dir / params.experiment_name / f'epoch_{epoch_i}.ch')

    test_fun = functools.partial(trainer.test, callbacks=[MAPCallback(list(RawPreprocessor.labels2id.keys())),
                                                          AccuracyCallback(), SaveBestCallback(params)])
    try:
        trainer.train(after_epoch_funcs=[save_last, save_each, test_fun])
    except KeyboardInterrupt:
        logger.error('Training process was interrupted.')
        trainer.save_state_dict(params.dump_dir / params.experiment_name / 'interrupt.ch')
    except Exception as e:
        logger.error(e)
        raise
    return trainer


def main(params, model_params):
    show_params(model_params, 'model')
    show_params(params, 'trainer')
    set_seed(params.seed)
    if not (torch.cuda.is_available() and params.gpu):
        raise RuntimeError('this entry point drives the sm_100a kernels and needs --gpu with a CUDA device; '
                           'the CPU path of the recipe is the reference itself')
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if world > 1:  # torchrun
        params.local_rank = int(os.environ['RANK'])
        local = int(os.environ.get('LOCAL_RANK', params.local_rank))
        torch.cuda.set_device(local)
        torch.distributed.init_process_group(backend=params.dist_backend)
    elif params.local_rank != -1 and params.dist_world_size > 1:  # the reference's explicit flags
        local = params.local_rank % torch.cuda.device_count()
        torch.cuda.set_device(local)
        torch.distributed.init_process_group(backend=params.dist_backend, init_method=params.dist_init_method,
                                             world_size=params.dist_world_size, rank=params.local_rank)
    else:
        local = torch.cuda.current_device()
        params.local_rank = -1
    params.distributed = torch.distributed.is_initialized()
    return run_worker(torch.device('cuda', local), params, model_params)


if __name__ == '__main__':
    import multiprocessing as mp
    (parser, model_parser), (params, model_params) = get_params((get_trainer_parser, get_model_parser))
    os.makedirs(params.dump_dir / params.experiment_name, exist_ok=True)
    params.log_file = params.dump_dir / params.experiment_name / f'{datetime.now().strftime("%d-%m-%Y_%H-%M-%S")}.log' \
        if params.local_rank in [-1, 0] else None
    params.n_jobs = min(params.n_jobs, mp.cpu_count() // 2)
    if params.local_rank in [0, -1]:
        write_config_file(parser, params, params.dump_dir / params.experiment_name / 'trainer.cfg')
        write_config_file(model_parser, model_params, params.dump_dir / params.experiment_name / 'model.cfg')
    main(params, model_params)
