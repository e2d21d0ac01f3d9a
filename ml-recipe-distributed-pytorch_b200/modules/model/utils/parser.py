"""Option surface of the recipe: same flag names, types, defaults and config-file behaviour as the
reference's modules/model/utils/parser.py:9-207, so that config/test_bert.cfg and
config/validate.cfg parse unchanged.  Declared as tables instead of call sequences; the one
deliberate extension is `--model bert-large-uncased` (BASELINE config 4 needs it; the
reference's choice list, parser.py:66-68, stops at bert-base-uncased / roberta-base).
"""
import logging
from pathlib import Path

try:
    import configargparse
except ImportError:  # this image has no configargparse: use the in-tree stand-in
    from .... import configargparse_shim as configargparse

logger = logging.getLogger(__name__)


def cast2(type_):
    """'None' on the command line / in a cfg means None (reference parser.py:34-35)."""
    return lambda x: type_(x) if x != 'None' else None


# (flags, kwargs) tables -----------------------------------------------------------------
_CFG = dict(required=False, is_config_file=True)
_FLAG = dict(action='store_true')

MODEL_OPTIONS = [
    (('-c', '--config_file'), dict(_CFG, help='Config file path.')),
    (('--model_config_file',), dict(_CFG, help='Model config file path.')),
    (('--model',), dict(type=str, default='bert-base-uncased',
                        choices=['bert-base-uncased', 'roberta-base', 'bert-large-uncased'],
                        help='Transformer model name.')),
    (('--hidden_dropout_prob',), dict(type=float, default=0.1, help='Model dropout probability.')),
    (('--attention_probs_dropout_prob',), dict(type=float, default=0.1, help='Attention dropout probability.')),
    (('--layer_norm_eps',), dict(type=float, default=1e-12, help='Layer norm eps.')),
    (('--vocab_file',), dict(type=cast2(str), default=None, help='Path to WordPiece/BPE vocab.')),
    (('--merges_file',), dict(type=cast2(str), default=None, help='BPE merge table path.')),
    (('--lowercase',), dict(_FLAG, help='Tokenize lowercase strings.')),
    (('--handle_chinese_chars',), dict(_FLAG, help='Do not replace chinese symbols with UNK tokens.')),
]

BASE_OPTIONS = [
    (('-c', '--config_file'), dict(_CFG, help='Config file path.')),
    (('--data_path',), dict(type=str, required=True, help='Path to JSON with documents.')),
    (('--processed_data_path',), dict(type=str, required=True, help='Path where processed dataset is saved.')),
    (('--gpu',), dict(_FLAG, help='Use gpu to train or validate models.')),
    (('--max_seq_len',), dict(type=int, default=384, help='Max input seq length.')),
    (('--max_question_len',), dict(type=int, default=64, help='Max question length.')),
    (('--doc_stride',), dict(type=int, default=128, help='Step size during doc splitting.')),
    (('--split_by_sentence',), dict(_FLAG, help='Split document by sentence instead.')),
    (('--truncate',), dict(_FLAG, help='Cut off long sentences during splitting by sentence.')),
    (('--n_jobs',), dict(type=int, default=16, help='Number of threads used in dataloader.')),
]

TRAINER_OPTIONS = [
    (('--trainer_config_file',), dict(_CFG, help='Trainer config file path.')),
    (('--dump_dir',), dict(type=Path, default='../results', help='Dump path.')),
    (('--experiment_name',), dict(type=str, required=True, help='Experiment name.')),
    (('--last',), dict(type=cast2(str), default=None, help='Restored checkpoint.')),
    (('--seed',), dict(type=cast2(int), default=None, help='Seed for random state.')),
    (('--n_epochs',), dict(type=int, default=10, help='Number of epochs.')),
    (('--train_batch_size',), dict(type=int, default=128, help='Number of items in batch.')),
    (('--test_batch_size',), dict(type=int, default=16, help='Number of items in batch.')),
    (('--batch_split',), dict(type=int, default=1, help='Batch is split into this number of chunks.')),
    (('--lr',), dict(type=float, default=1e-5, help='Learning rate for optimizer.')),
    (('--weight_decay',), dict(type=float, default=0.01, help='Weight decay for optimizer.')),
    (('--clear_processed',), dict(_FLAG, help='Clear previous processed dataset.')),
    (('--w_start',), dict(type=float, default=1, help='Weight of start position classification.')),
    (('--w_end',), dict(type=float, default=1, help='Weight of end position classification.')),
    (('--w_start_reg',), dict(type=float, default=0, help='Weight of start position regression loss.')),
    (('--w_end_reg',), dict(type=float, default=0, help='Weight of end position regression loss.')),
    (('--w_cls',), dict(type=float, default=1, help='Weight of doc label classification.')),
    (('--loss',), dict(type=str, default='ce', choices=['ce', 'focal', 'smooth'],
                       help='Type of doc label classification loss')),
    (('--smooth_alpha',), dict(type=float, default=0.01, help='Smooth CE loss parameter.')),
    (('--focal_alpha',), dict(type=float, default=1, help='Focal loss parameter.')),
    (('--focal_gamma',), dict(type=float, default=2, help='Focal loss parameter.')),
    (('--max_grad_norm',), dict(type=float, default=1, help='Max norm of the gradients')),
    (('--sync_bn',), dict(_FLAG, help='Synchronize batch norm parameters (no-op: BERT has none).')),
    (('--warmup_coef',), dict(type=float, default=0.05, help='Warmup coefficient.')),
    (('--apex_level',), dict(type=cast2(str), choices=[None, 'O0', 'O1', 'O2', 'O3'], default=None,
                             help='Accepted for compatibility: compute is always bf16 with fp32 master weights.')),
    (('--apex_verbosity',), dict(type=int, default=1, help='Accepted for compatibility.')),
    (('--apex_loss_scale',), dict(type=cast2(float), default=None, help='Accepted for compatibility.')),
    (('--drop_optimizer',), dict(_FLAG, help='Not restore optimizer and scheduler from checkpoint.')),
    (('--debug',), dict(_FLAG, help='Debug mode.')),
    (('--dummy_dataset',), dict(_FLAG, help='Use generated dataset instead real data.')),
    (('--local_rank',), dict(type=int, default=-1, help='Local rank of process during distributed training.')),
    (('--dist_backend',), dict(type=str, default='nccl', choices=['nccl'], help='Distributed training backend.')),
    (('--dist_init_method',), dict(type=str, default='tcp://127.0.0.1:9080', help='Distributed init method.')),
    (('--dist_world_size',), dict(type=int, default=1, help='Number of machines used during training.')),
    (('--best_metric',), dict(choices=['map'], type=str, default='map', help='Best metric name.')),
    (('--best_order',), dict(choices=['>', '<'], type=str, default='>', help='Best metric order.')),
    (('--finetune',), dict(_FLAG, help='Turn on finetune mode.')),
    (('--finetune_transformer',), dict(_FLAG, help='Finetune transformer module.')),
    (('--finetune_position',), dict(_FLAG, help='Finetune classification head.')),
    (('--finetune_position_reg',), dict(_FLAG, help='Finetune regression head.')),
    (('--finetune_class',), dict(_FLAG, help='Finetune doc label classification head.')),
    (('--bpe_dropout',), dict(type=cast2(float), default=None, help='Use BPE dropout.')),
    (('--optimizer',), dict(type=str, default='adam', choices=['adam', 'adamod'], help='Optimizer name.')),
    (('--train_label_weights',), dict(_FLAG, help='Use label weights in CE loss.')),
    (('--train_sampler_weights',), dict(_FLAG, help='Use oversampling.')),
    (('--log_file',), dict(type=str, default=None, help='Ignored on input; holds the log path after dump.')),
]

PREDICTOR_OPTIONS = [
    (('--predictor_config_file',), dict(_CFG, help='Predictor config file path.')),
    (('--checkpoint',), dict(required=True, type=cast2(str), help='Restored checkpoint path.')),
    (('--batch_size',), dict(type=int, default=16, help='Batch size.')),
    (('--buffer_size',), dict(type=int, default=4096, help='Buffer queue size.')),
    (('--limit',), dict(type=cast2(int), default=None, help='Process only specified number of documents.')),
]


def _build(description, *tables):
    parser = configargparse.ArgumentParser(description=description)
    for table in tables:
        for flags, kwargs in table:
            parser.add_argument(*flags, **kwargs)
    return parser


def get_model_parser():
    return _build('Model config parser.', MODEL_OPTIONS)


def get_trainer_parser():
    return _build('Trainer config parser.', BASE_OPTIONS, TRAINER_OPTIONS)


def get_predictor_parser():
    return _build('Validation config parser.', BASE_OPTIONS, PREDICTOR_OPTIONS)


def get_params(parser_getters, args=None):
    """Every parser takes what it knows; a flag unknown to ALL parsers aborts (parser.py:9-31)."""
    parsers, params, unused = [], [], None
    for getter in parser_getters:
        parser = getter()
        known, rest = parser.parse_known_args(args)
        parsers.append(parser)
        params.append(known)
        rest = set(rest)
        unused = rest if unused is None else unused & rest
    if unused:
        for parser in parsers:
            parser.print_help()
        print(f'Incorrect command line parameters: {unused}.')
        raise SystemExit(0)
    return parsers, params


def write_config_file(parser, parsed_namespace, output_path):
    items = {k: getattr(parsed_namespace, k) for k in sorted(vars(parsed_namespace)) if 'config' not in k}
    try:
        with open(output_path, 'w') as f:
            f.write(parser._config_file_parser.serialize(items))
    except IOError:
        logger.error(f'Could not open file {output_path}.')
        raise
    logger.info(f'Config was saved to {output_path}.')


def load_config_file(parser_getter, config_path):
    parser = parser_getter()
    return parser, parser.parse_args(f'-c {config_path}')
