"""Host-side logic that needs no GPU: option surface, collate / dummy data ABI, bucket plan and the
bucketed gradient all-reduce under gloo (world_size 2), dropout-hash determinism, schedules."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mods(pkg):
    imp = lambda n: importlib.import_module("bqa_b200.modules." + n)  # noqa: E731
    return dict(parser=imp("model.utils.parser"), dataset=imp("model.dataset"), init=imp("init"),
                reducer=imp("model.trainer.reducer"), trainer=imp("model.trainer.trainer"),
                optim=imp("model.trainer.optim"), model=imp("model.model"))


CFG_DIRS = [os.path.join(ROOT, "config")] + (["/root/reference/config"] if os.path.isdir("/root/reference/config") else [])


@pytest.mark.parametrize("cfg_dir", CFG_DIRS)
def test_trainer_cfg_parses_unchanged(mods, cfg_dir):
    P = mods["parser"]
    argv = ["-c", os.path.join(cfg_dir, "test_bert.cfg"), "--max_seq_len", "384", "--train_batch_size", "8",
            "--batch_split", "1", "--apex_level", "None"]
    (_, _), (t, m) = P.get_params((P.get_trainer_parser, P.get_model_parser), argv)
    assert (t.max_seq_len, t.train_batch_size, t.batch_split, t.apex_level) == (384, 8, 1, None)  # CLI beats cfg
    assert t.debug is True and t.dummy_dataset is True and t.sync_bn is True and t.finetune is False
    assert t.loss == "smooth" and t.smooth_alpha == 0.01 and t.warmup_coef == 0.6 and t.lr == 1e-5
    assert t.weight_decay == 1e-4 and t.seed is None and t.last is None and t.n_epochs == 2
    assert m.model == "bert-base-uncased" and m.lowercase is True and m.handle_chinese_chars is False
    assert m.hidden_dropout_prob == 0.1 and m.layer_norm_eps == 1e-12 and m.merges_file is None


@pytest.mark.parametrize("cfg_dir", CFG_DIRS)
def test_predictor_cfg_parses_unchanged(mods, cfg_dir):
    P = mods["parser"]
    (_, _), (p, m) = P.get_params((P.get_predictor_parser, P.get_model_parser), ["-c", os.path.join(cfg_dir, "validate.cfg")])
    assert (p.batch_size, p.limit, p.max_seq_len, p.doc_stride, p.buffer_size) == (16, 100, 512, 128, 4096)
    assert p.gpu is True and p.split_by_sentence is True


def test_unknown_flag_aborts_and_config_roundtrip(mods, tmp_path):
    P = mods["parser"]
    with pytest.raises(SystemExit):
        P.get_params((P.get_trainer_parser, P.get_model_parser),
                     ["-c", os.path.join(ROOT, "config", "test_bert.cfg"), "--no_such_flag", "1"])
    (tp, _), (t, _) = P.get_params((P.get_trainer_parser, P.get_model_parser),
                                   ["-c", os.path.join(ROOT, "config", "test_bert.cfg")])
    out = tmp_path / "trainer.cfg"
    P.write_config_file(tp, t, out)
    text = out.read_text()
    assert "max_seq_len = 512" in text and "config_file" not in text


def test_collate_and_dummy_dataset_abi(mods):
    D, I = mods["dataset"], mods["init"]
    tok = I.DummyTokenizer("bert", 30522)
    np.random.seed(1234)
    ds = D.DummyDataset(tok, max_seq_len=96, max_question_len=16)
    items = [ds[i] for i in range(3)]
    inputs, labels = D.collate_fun(items, tok)
    ids = inputs["input_ids"]
    assert ids.dtype == torch.int64 and ids.shape == (3, 96)
    assert inputs["attention_mask"].dtype == torch.bool and inputs["attention_mask"].all()
    assert (ids[:, 0] == 101).all() and (ids[:, 17] == 102).all() and (ids[:, -1] == 102).all()
    body = torch.cat([ids[:, 1:17], ids[:, 18:-1]], 1)
    assert not ((body == 0) | (body == 101) | (body == 102)).any()
    assert (inputs["token_type_ids"][:, :18] == 0).all() and (inputs["token_type_ids"][:, 18:] == 1).all()
    assert labels["start_class"].tolist() == [0, 0, 0] and labels["end_class"].tolist() == [95] * 3
    assert labels["start_reg"].dtype == torch.float32 and labels["cls"].dtype == torch.int64
    # ragged batch: padded to the batch maximum with mask = ids > 0
    items[1].input_ids = items[1].input_ids[:40] + [102]
    inputs, _ = D.collate_fun(items, tok)
    assert inputs["attention_mask"][1].sum() == 41 and (inputs["input_ids"][1, 41:] == 0).all()
    assert (inputs["token_type_ids"][1, 41:] == 1).all()


def test_collate_matches_reference_collate(mods):
    if not os.path.isdir("/root/reference/modules"):
        pytest.skip("reference not present")
    from oracle.reference_shims import StubTokenizer, import_reference
    ref = import_reference()
    D = mods["dataset"]
    tok = StubTokenizer()
    rng = np.random.RandomState(0)
    items, ref_items = [], []
    for i, n in enumerate((30, 57, 12)):
        ids = [101] + rng.randint(1000, 2000, 7).tolist() + [102] + rng.randint(1000, 2000, n).tolist() + [102]
        kw = dict(example_id=str(i), input_ids=ids, start_id=i, end_id=i + 3, label_id=i, start_position=0.1 * i,
                  end_position=0.2 * i)
        items.append(D.DatasetItem(**kw))
        ref_items.append(ref.dataset.DatasetItem(**kw))
    a_in, a_lab = D.collate_fun(items, tok)
    b_in, b_lab = ref.dataset.collate_fun(ref_items, tok)
    for k in b_in:
        assert a_in[k].dtype == b_in[k].dtype and torch.equal(a_in[k], b_in[k]), k
    for k in b_lab:
        assert a_lab[k].dtype == b_lab[k].dtype and torch.equal(a_lab[k], b_lab[k]), k


def test_bucket_plan(mods):
    R = mods["reducer"]
    begins = [900, 700, 500, 300, 0]
    plan = R.bucket_plan(begins, 1000, min_bucket_elems=250)
    assert plan == [(1, 700, 1000), (3, 300, 700), (4, 0, 300)]
    assert R.bucket_plan(begins, 1000, min_bucket_elems=1) == [(0, 900, 1000), (1, 700, 900), (2, 500, 700),
                                                                (3, 300, 500), (4, 0, 300)]
    assert R.bucket_plan(begins, 1000, min_bucket_elems=10 ** 9) == [(4, 0, 1000)]
    covered = sorted((b, e) for _, b, e in plan)
    assert covered[0][0] == 0 and covered[-1][1] == 1000
    assert all(covered[i][1] == covered[i + 1][0] for i in range(len(covered) - 1))


def _gloo_worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from __graft_entry__ import load_package
    load_package()
    R = importlib.import_module("bqa_b200.modules.model.trainer.reducer")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    torch.manual_seed(rank)
    flat = torch.randn(1000)
    mine = flat.clone()
    red = R.GradReducer(flat, [900, 700, 500, 300, 0], min_bucket_mb=250 * 4 / (1 << 20))
    assert [p[:2] for p in red.phases()] == [(0, 1), (2, 3), (4, 4)]
    for first, last, bucket in red.phases():
        red.reduce_bucket(bucket)
    red.wait()
    torch.save((mine, flat), os.path.join(tmp, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gradient_allreduce_mean_gloo(tmp_path, world):
    import torch.multiprocessing as mp
    port = 29500 + os.getpid() % 2000 + 7 * world
    mp.spawn(_gloo_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    loaded = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    mean = sum(m for m, _ in loaded) / world
    for _, f in loaded:
        torch.testing.assert_close(f, mean)          # reference semantic: DDP averages over ranks
        assert torch.equal(f, loaded[0][1])          # and every rank holds the same result


def test_default_bucket_plan_is_two_buckets_for_bert_base(pkg):
    """reducer.DEFAULT_BUCKET_MB (220 MB): heads + layers 11..4 | layers 3..0 + embeddings for BERT-base-sized phases
    (every overlapped all-reduce costs 0.1-0.2 ms whatever its size)."""
    R = importlib.import_module("bqa_b200.modules.model.trainer.reducer")
    mb = lambda x: int(x * (1 << 20) / 4)
    sizes = [mb(2.4)] + [mb(28.3)] * 12 + [mb(95.0)]   # heads, layers 11..0, embeddings (elements)
    total = sum(sizes)
    begins, left = [], total
    for n in sizes:
        left -= n
        begins.append(left)
    plan = R.bucket_plan(begins, total, mb(R.DEFAULT_BUCKET_MB))
    assert len(plan) == 2 and plan[0][0] == 8 and plan[1][0] == 13          # last phase of each bucket
    assert plan[0][2] == total and plan[1][1] == 0 and plan[0][1] == plan[1][2]
    per_layer = R.bucket_plan(begins, total, mb(25.0))
    assert len(per_layer) == 13
    # the bf16 exchange has no CPU path: it must say so instead of silently exchanging fp32
    red = R.GradReducer(torch.zeros(16), [8, 0], world=2, exchange_dtype=torch.bfloat16, min_bucket_mb=0.0)
    with pytest.raises(RuntimeError, match="no CPU path"):
        red._all_reduce_bf16(red.flat[:8])


def _shard_worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from __graft_entry__ import load_package
    load_package()
    Pm = importlib.import_module("bqa_b200.modules.model.inference.predictor")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    docs = [[f"doc{d}-chunk{c}" for c in range(1 + d % 3)] for d in range(7)]
    shard = Pm.DocumentShard(docs, rank, world)
    mine = [shard[i] for i in range(len(shard))]
    # a Predictor-shaped holder of this rank's per-document winners (no model needed for the merge logic)
    pred = Pm.Predictor.__new__(Pm.Predictor)
    from collections import defaultdict
    pred.scores, pred.candidates, pred.items = defaultdict(int), {}, {}
    for chunks in mine:
        doc = chunks[0].split("-")[0]
        pred.scores[doc] = float(len(chunks)) + rank
        pred.candidates[doc] = Pm.PredictorCandidate(start_id=len(chunks), end_id=len(chunks) + 1, start_reg=0.5,
                                                     end_reg=0.75, label=rank)
        pred.items[doc] = chunks[-1]
    n = Pm.gather_predictions(pred)
    torch.save((mine, n, dict(pred.scores), {k: (v.start_id, v.label) for k, v in pred.candidates.items()},
                dict(pred.items)), os.path.join(tmp, f"s{rank}.pt"))
    dist.destroy_process_group()


def test_document_sharded_prediction_merge_gloo_world2(tmp_path):
    """BASELINE configs[4] on N ranks: documents (not chunks) are dealt to the ranks, every rank keeps the best
    candidate of ITS documents (predictor.py:63-87), rank 0 ends up with all of them."""
    import torch.multiprocessing as mp
    port = 31500 + os.getpid() % 2000
    mp.spawn(_shard_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    (d0, n0, sc0, c0, it0), (d1, n1, sc1, c1, it1) = (torch.load(tmp_path / f"s{r}.pt") for r in range(2))
    all_docs = sorted(ch[0].split("-")[0] for ch in d0 + d1)
    assert all_docs == [f"doc{d}" for d in range(7)]                          # disjoint and complete
    assert all(len({c.split("-")[0] for c in chunks}) == 1 for chunks in d0 + d1)  # a document never straddles ranks
    assert n0 == 7 and n1 == 0 and sorted(c0) == all_docs
    assert c0["doc1"] == (2, 1) and c0["doc2"] == (3, 0) and sc0["doc1"] == 3.0  # rank 1's winner reached rank 0
    assert it0["doc4"] == "doc4-chunk1" and it0["doc5"] == "doc5-chunk2"
    assert sorted(c1) == ["doc1", "doc3", "doc5"]                             # other ranks keep only their own


def test_dropout_hash_is_deterministic_and_calibrated():
    from oracle.dropout_hash import drop_key, keep_mask
    a = keep_mask(1234, 7, 0.1, 200000)
    b = keep_mask(1234, 7, 0.1, 200000)
    c = keep_mask(1234, 8, 0.1, 200000)
    assert torch.equal(a, b) and not torch.equal(a, c)
    assert abs(a.float().mean().item() - 0.9) < 3e-3
    assert torch.equal(keep_mask(1234, 7, 0.1, 1000, start=500), a[500:1500])
    assert keep_mask(1, 1, 0.0, 10).all()
    assert drop_key(1234, 7) != drop_key(1234, 8)


def test_warmup_lambda_matches_oracle(mods):
    from oracle.bertqa_oracle import linear_schedule_lambda
    fn = mods["trainer"].linear_warmup_lambda(6, 20)
    assert [fn(t) for t in range(25)] == [linear_schedule_lambda(t, 6, 20) for t in range(25)]


def test_unsupported_options_fail_loudly(mods):
    import types
    I = mods["init"]
    with pytest.raises(NotImplementedError):
        I.init_loss(types.SimpleNamespace(loss="hinge", smooth_alpha=0.01), {})
    mp = types.SimpleNamespace(model="xlnet-base-cased", vocab_file=None)
    with pytest.raises(NotImplementedError):
        I.init_model(mp)
    # both of the reference's --model choices are known (model.py:9-10); RoBERTa's special ids and conventions
    tok = I.DummyTokenizer("roberta", 50265)
    assert (tok.cls_token_id, tok.pad_token_id, tok.sep_token_id, tok.unk_token_id) == (0, 1, 2, 3)
    shape = importlib.import_module("bqa_b200.engine").MODEL_SHAPES["roberta-base"]
    assert (shape["vocab"], shape["max_pos"], shape["type_vocab"], shape["pad_token_id"], shape["position_mode"]) == \
        (50265, 514, 1, 1, 1)
    # loss=focal and optimizer=adamod are configuration holders on the host (kernels: tests -m gpu)
    wl = I.init_loss(types.SimpleNamespace(loss="focal", focal_alpha=0.5, focal_gamma=3, smooth_alpha=0.01), {})
    assert type(wl._losses['cls'][0]).__name__ == "FocalLossWithLogits"
    with pytest.raises(NotImplementedError):
        mods["model"].FocalLossWithLogits(ignore_index=-100)
    with pytest.raises(ValueError, match="beta3"):
        mods["optim"].AdaMod([torch.nn.Parameter(torch.zeros(2))], None, beta3=1.5)


def test_model_skeleton_has_reference_state_dict_keys(mods):
    import types
    from oracle import bertqa_oracle as O
    shape = dict(num_layers=2, hidden=64, heads=1, intermediate=128, vocab=200, max_pos=128, type_vocab=2)
    mp = types.SimpleNamespace(model="bert-base-uncased", model_shape=shape, hidden_dropout_prob=0.1,
                               attention_probs_dropout_prob=0.1, layer_norm_eps=1e-12)
    m = mods["model"].BertForQuestionAnswering(mp, 5)
    names = [k for k, _ in m.named_parameters()]
    cfg = O.OracleConfig(num_layers=2, hidden=64, heads=1, intermediate=128, vocab=200, max_pos=128)
    assert names == O.param_names(cfg) == list(m.state_dict().keys())
    shapes = O.param_shapes(cfg)
    assert all(tuple(p.shape) == tuple(shapes[k]) for k, p in m.named_parameters())
    for attr in ("transformer", "position_outputs", "classifier", "reg_start", "reg_end", "model_params"):
        assert hasattr(m, attr)
    assert m.transformer.config.hidden_size == 64 and m.transformer.config.num_labels == 5
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 8, dtype=torch.int64))


def test_compact_collate_matches_reference_layout_and_raises_without_sep():
    """collate_fun(compact=True) ships the same ids as int32 and keeps the reference's ValueError
    (row.index(sep_token_id), split_dataset.py:495) for a row without [SEP]."""
    import importlib
    import types
    ds = importlib.import_module("bqa_b200.modules.model.dataset")
    tok = types.SimpleNamespace(pad_token_id=0, sep_token_id=102, cls_token_id=101, model_name='bert')
    items = [ds.DatasetItem(example_id='a', input_ids=[101, 5, 6, 102, 7, 8, 9, 102], start_id=4, end_id=6, label_id=1,
                            start_position=0.5, end_position=0.75),
             ds.DatasetItem(example_id='b', input_ids=[101, 5, 102, 7, 102], start_id=-1, end_id=-1, label_id=0,
                            start_position=0.0, end_position=1.0)]
    full, labels = ds.collate_fun(items, tok)
    compact, labels_c = ds.collate_fun(items, tok, compact=True)
    assert compact['input_ids_i32'].dtype == torch.int32
    assert torch.equal(compact['input_ids_i32'].long(), full['input_ids'])
    assert compact['sep_token_id'] == 102 and compact['type_coef'] == 1
    for k in labels:
        assert torch.equal(labels[k], labels_c[k])
    bad = [ds.DatasetItem(example_id='c', input_ids=[101, 5, 6], start_id=0, end_id=0, label_id=0,
                          start_position=0.0, end_position=1.0)]
    with pytest.raises(ValueError):
        ds.collate_fun(bad, tok, compact=True)
