"""Logger / seeding / parameter dump helpers (reference modules/utils.py:10-51)."""
import logging
import random

import numpy as np
import torch

logger = logging.getLogger(__name__)


def get_logger(*, level=logging.INFO, filename=None, filemode='w', logger_name=None, debug=False):
    for handler in logging.root.handlers[:]:
        logging.root.removeHandler(handler)
    handlers = [logging.StreamHandler()]
    if filename is not None:
        handlers.append(logging.FileHandler(filename, filemode))
    where = '%(pathname)s:%(funcName)s:%(lineno)d' if debug else '%(name)s'
    logging.basicConfig(format=f'%(asctime)s - %(levelname)s - {where} -   %(message)s', datefmt='%m/%d/%Y %H:%M:%S',
                        level=level, handlers=handlers)
    log = logging.getLogger(__file__ if logger_name is None else logger_name)
    if filename is not None and filemode == 'w':
        log.info(f'All logs will be dumped to {filename}.')
    return log


def set_seed(seed=None):
    if seed is None:
        return
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    logger.info(f'Random seed was set to {seed}.')


def show_params(params, name):
    logger.info(f'Input {name} parameters:')
    for k in sorted(vars(params)):
        logger.info(f'\t\t{k}: {getattr(params, k)}')
