"""CPU: host-side mirror of the reference's module interface (no kernels run)."""
import os
import sys

import numpy as np
import pytest
import torch

import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200.modules.quantize import kernel_distance, kernel_mode
from rq_vae_recommender_b200 import functional as RF
from conftest import ROOT, load_golden


def test_enum_values_match_reference():
    assert [m.value for m in R.QuantizeForwardMode] == [1, 2, 3]
    assert R.QuantizeForwardMode.ROTATION_TRICK.name == "ROTATION_TRICK"
    assert [d.value for d in R.QuantizeDistance] == [1, 2]
    assert kernel_mode(R.QuantizeForwardMode.STE, training=False) == RF.MODE_EVAL
    assert kernel_mode(R.QuantizeForwardMode.GUMBEL_SOFTMAX, True) == RF.MODE_GUMBEL_SOFTMAX
    with pytest.raises(Exception, match="Unsupported Quantize forward mode"):
        kernel_mode("nope", True)
    with pytest.raises(Exception, match="Unsupported Quantize distance mode"):
        kernel_distance("nope")


def test_state_dict_keys_and_load_from_reference_fixture():
    z = load_golden("rqvae_mlp_small.npz")
    sd = {k[4:]: torch.from_numpy(z[k]) for k in z.keys() if k.startswith("sd::")}
    m = R.RqVae(input_dim=64, embed_dim=32, hidden_dims=[48, 40], codebook_size=64, codebook_kmeans_init=False,
                codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=3, n_cat_features=0)
    assert sorted(m.state_dict().keys()) == sorted(sd.keys())
    m.load_state_dict(sd)                     # a reference checkpoint's "model" entry loads unchanged
    assert "kmeans_initted" not in m.state_dict()
    q = R.Quantize(16, 8, sim_vq=True, codebook_normalize=True)
    assert sorted(q.state_dict().keys()) == ["embedding.weight", "out_proj.0.weight"]
    assert float(q.weight.min()) >= 0.0 and float(q.weight.max()) <= 1.0      # U[0,1) init


def test_reference_checkpoint_file_loads_without_the_reference_tree(tmp_path, capsys):
    """tests/golden/rqvae_checkpoint_ref.pt was written by the reference exactly as train_rqvae.py:209-221 does
    ({"iter", "model", "model_config", "optimizer"}); its `model_config` pickles the reference's own classes, which do
    not exist here.  `load_pretrained` (rqvae.py:103-106) must still restore every tensor bit for bit."""
    import os
    from conftest import GOLDEN
    z = load_golden("rqvae_checkpoint_ref.npz")
    m = R.RqVae(input_dim=24, embed_dim=8, hidden_dims=[16], codebook_size=16, codebook_kmeans_init=False,
                codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=3, n_cat_features=0)
    m.load_pretrained(os.path.join(GOLDEN, "rqvae_checkpoint_ref.pt"))
    assert "Loaded RQVAE Iter 41" in capsys.readouterr().out
    sd = m.state_dict()
    want = {k[4:]: torch.from_numpy(z[k]) for k in z.keys() if k.startswith("sd::")}
    assert sorted(sd.keys()) == sorted(want.keys())
    for k in want:
        assert torch.equal(sd[k].cpu(), want[k]), k
    # and the way back: a checkpoint written with the same four keys from the mirror loads into a fresh mirror
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3)
    state = {"iter": 7, "model": m.state_dict(), "model_config": m.config, "optimizer": opt.state_dict()}
    path = tmp_path / "checkpoint_7.pt"
    torch.save(state, str(path))
    m2 = R.RqVae(input_dim=24, embed_dim=8, hidden_dims=[16], codebook_size=16, codebook_kmeans_init=False,
                 codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=3, n_cat_features=0)
    m2.load_pretrained(str(path))
    for k in want:
        assert torch.equal(m2.state_dict()[k].cpu(), want[k]), k


def test_checkpoint_round_trip_with_the_live_reference(tmp_path):
    """Both directions against the UNMODIFIED reference (container only): the reference loads a checkpoint the mirror
    wrote, the mirror loads one the reference wrote."""
    import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not mounted")
    ref = ref_import.load()
    kw = dict(input_dim=24, embed_dim=8, hidden_dims=[16], codebook_size=16, codebook_kmeans_init=False, n_layers=3,
              n_cat_features=0)
    mine = R.RqVae(codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, **kw)
    theirs = ref.rqvae.RqVae(codebook_mode=ref.quantize.QuantizeForwardMode.ROTATION_TRICK, **kw)
    p1 = tmp_path / "from_mirror.pt"
    torch.save({"iter": 3, "model": mine.state_dict(), "model_config": {k: v for k, v in mine.config.items()
                                                                         if k not in ("self", "__class__")},
                "optimizer": torch.optim.AdamW(mine.parameters()).state_dict()}, str(p1))
    theirs.load_pretrained(str(p1))
    for k, v in mine.state_dict().items():
        assert torch.equal(theirs.state_dict()[k], v), k
    with torch.no_grad():
        for prm in theirs.parameters():
            prm.add_(0.5)
    p2 = tmp_path / "from_reference.pt"
    torch.save({"iter": 4, "model": theirs.state_dict(), "model_config": theirs.config,
                "optimizer": torch.optim.AdamW(theirs.parameters()).state_dict()}, str(p2))
    mine.load_pretrained(str(p2))
    for k, v in theirs.state_dict().items():
        assert torch.equal(mine.state_dict()[k], v), k


def test_constructor_signatures_match_reference():
    import inspect
    sig = inspect.signature(R.Quantize.__init__)
    assert list(sig.parameters)[1:] == ["embed_dim", "n_embed", "do_kmeans_init", "codebook_normalize", "sim_vq",
                                        "commitment_weight", "forward_mode", "distance_mode"]
    assert sig.parameters["forward_mode"].default == R.QuantizeForwardMode.GUMBEL_SOFTMAX
    sig = inspect.signature(R.RqVae.__init__)
    assert list(sig.parameters)[1:] == ["input_dim", "embed_dim", "hidden_dims", "codebook_size",
                                        "codebook_kmeans_init", "codebook_normalize", "codebook_sim_vq",
                                        "codebook_mode", "n_layers", "commitment_weight", "n_cat_features"]
    sig = inspect.signature(R.SemanticIdTokenizer.__init__)
    assert list(sig.parameters)[1:] == ["input_dim", "output_dim", "hidden_dims", "codebook_size", "n_layers",
                                        "n_cat_feats", "commitment_weight", "rqvae_weights_path",
                                        "rqvae_codebook_normalize", "rqvae_sim_vq"]
    sig = inspect.signature(R.Kmeans.__init__)
    assert list(sig.parameters)[1:4] == ["k", "max_iters", "stop_threshold"]
    assert sig.parameters["stop_threshold"].default == 1e-10


def test_cpu_tensors_are_rejected_not_silently_computed():
    q = R.Quantize(32, 16, do_kmeans_init=False, forward_mode=R.QuantizeForwardMode.STE)
    with pytest.raises(R.RqEngineError, match="no CPU"):
        q(torch.randn(5, 32), 0.2)
    with pytest.raises(AssertionError):
        q(torch.randn(5, 31), 0.2)            # quantize.py:100
    tok = R.SemanticIdTokenizer(24, 8, [16], 6, n_cat_feats=0)
    assert tok.sem_ids_dim == 4 and not tok.rq_vae.training      # semids.py:58
    with pytest.raises(Exception, match="No match can be found in empty cache."):
        tok.exists_prefix(torch.zeros(4, 2, dtype=torch.int64))
    assert not tok.rq_vae.layers[0].do_kmeans_init


def test_product_never_imports_oracle():
    """The oracle is test infrastructure; the shipped package must not reference it."""
    pkg = os.path.join(ROOT, "rq_vae_recommender_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "rq_oracle" not in src and "import oracle" not in src and "from oracle" not in src, f


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from rq_vae_recommender_b200.init.kmeans import Kmeans
        from rq_vae_recommender_b200.parallel import shard_range
        g = torch.Generator().manual_seed(0)
        full = torch.randn(11, 4, generator=g)
        lo, hi = shard_range(11, rank, world)
        x = full[lo:hi]
        assert not Kmeans(k=3)._is_dist()        # rank-local unless asked: no collective behind the reference's call
        km = Kmeans(k=3, distributed=True)
        off, n = km._shard_layout(x.shape[0], x.device)
        rows = km._global_rows(x, [0, 5, 10, 6], off)
        np.random.seed(123 + rank)               # ranks disagree on purpose: rank 0's draw must win
        km._init_centroids(x)
        # corpus tokenisation's one collective: [N/G, L] id shards -> the full [N, L] table on every rank
        from rq_vae_recommender_b200.parallel import gather_row_shards
        ids_full = torch.arange(11 * 3, dtype=torch.int64).reshape(11, 3)
        gathered = gather_row_shards(ids_full[lo:hi].contiguous(), 11)
        # dedup column of the sharded corpus table: local rank within the key + counts of the lower ranks
        from rq_vae_recommender_b200.parallel import sharded_corpus_table

        def cpu_table(ids, _k):                  # stable cumcount in corpus order (semids.py:86-99), CPU stand-in
            seen, col = {}, []
            for row in ids.tolist():
                col.append(seen.get(tuple(row), 0))
                seen[tuple(row)] = col[-1] + 1
            return torch.cat([ids, torch.tensor(col, dtype=torch.int64).reshape(-1, 1)], dim=1)
        gd = torch.Generator().manual_seed(5)
        corpus = torch.randint(0, 3, (23, 2), generator=gd)          # 9 keys, 23 items: plenty of duplicates
        a, b = shard_range(23, rank, world)
        dense = sharded_corpus_table(corpus[a:b].contiguous(), 23, 3, table_fn=cpu_table)
        gathered_path = sharded_corpus_table(corpus[a:b].contiguous(), 23, 3, max_dense_keys=1, table_fn=cpu_table)
        table_ok = torch.equal(dense, cpu_table(corpus, 3)) and torch.equal(gathered_path, cpu_table(corpus, 3))
        # codebook-gradient all-reduce (mean over ranks), the training step's one collective
        from rq_vae_recommender_b200.parallel import allreduce_codebook_grads

        class _Layer:
            def __init__(self, v):
                self.embedding = torch.nn.Embedding(5, 4)
                self.embedding.weight.grad = torch.full((5, 4), float(v))

        class _Model:
            layers = [_Layer(rank + 1), _Layer(10 * (rank + 1))]
        allreduce_codebook_grads(_Model)
        grads_ok = (torch.allclose(_Model.layers[0].embedding.weight.grad, torch.full((5, 4), 1.5))
                    and torch.allclose(_Model.layers[1].embedding.weight.grad, torch.full((5, 4), 15.0)))
        q.put((rank, off, n, torch.equal(rows, full[[0, 5, 10, 6]]), km.centroids.clone(),
               torch.equal(gathered, ids_full), grads_ok, table_ok))
    finally:
        dist.destroy_process_group()


def test_kmeans_sharding_helpers_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 500)
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    [p.join(60) for p in procs]
    assert res[0][1] == 0 and res[1][1] == 6 and res[0][2] == res[1][2] == 11
    assert res[0][3] and res[1][3]
    assert torch.equal(res[0][4], res[1][4])     # identical initial centroids on both ranks
    assert res[0][5] and res[1][5]               # row-shard all-gather reassembles the global order
    assert res[0][6] and res[1][6]               # codebook gradients averaged over the two ranks
    assert res[0][7] and res[1][7]               # sharded dedup column == single-process table, both strategies


def test_data_utils_batch_to_and_cycle():
    """data/utils.py:4-17 of the reference: cycle restarts the loader, batch_to keeps the batch type."""
    from rq_vae_recommender_b200.data import utils as U
    from rq_vae_recommender_b200.data.schemas import SeqBatch
    it = U.cycle([1, 2, 3])
    assert [next(it) for _ in range(7)] == [1, 2, 3, 1, 2, 3, 1]
    b = SeqBatch(*[torch.arange(4) + i for i in range(len(SeqBatch._fields))])
    out = U.batch_to(b, "cpu")
    assert type(out) is SeqBatch and all(torch.equal(a, c) for a, c in zip(out, b))
    assert torch.equal(U.batch_to(torch.ones(3), "cpu"), torch.ones(3))
    d = U.batch_to({"a": torch.zeros(2), "n": 5}, "cpu")
    assert d["n"] == 5 and torch.equal(d["a"], torch.zeros(2))
    with pytest.raises(TypeError):
        U.batch_to(3.5, "cpu")
    with pytest.raises(RuntimeError):
        U.DevicePrefetcher([torch.ones(1)], "cpu")


def test_bench_reference_arm_prints_exactly_one_json_line():
    """bench.py's contract: stdout carries ONE JSON line (library banners are sent to stderr; the line is written to
    the saved descriptor, so an os._exit on the multi-rank path cannot lose it).  The reference arm runs on CPU."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                        "--cpu-sample-items", "2048"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "items/s" and d["value"] > 0
    # "reference" when oracle/_ref holds the vendored reference (oracle/build_ref.py), else the oracle port
    have_ref = os.path.exists(os.path.join(root, "oracle", "_ref", "MANIFEST.json"))
    assert d["cpu_baseline"]["kind"] == ("reference" if have_ref else "port") and d["e2e"]["h2d_bytes_per_step"] == 0
    assert d["steps"] == 1 and d["warmup"] == 1          # the arm honours the driver's --steps / --warmup
    # a non-zero rank of a multi-rank launch does no work and prints nothing
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_mlp_engine_plan_is_cuda_inference_only():
    """The encoder-layer kernel is taken only for CUDA fp32 inference with TF32 matmuls allowed; on CPU (and under
    autograd) the MLP is the reference's nn.Sequential (modules/encoder.py:17-36), bit for bit."""
    import torch
    from rq_vae_recommender_b200.modules.encoder import MLP
    torch.manual_seed(0)
    mlp = MLP(48, [32, 16], 8, normalize=True).eval()
    x = torch.randn(5, 48)
    prev = torch.get_float32_matmul_precision()
    try:
        torch.set_float32_matmul_precision("high")
        with torch.no_grad():
            assert mlp._engine_plan(x) is None
            y = mlp(x)
            assert torch.equal(y, mlp.mlp(x))
    finally:
        torch.set_float32_matmul_precision(prev)
    assert [k for k in mlp.state_dict()] == ["mlp.0.weight", "mlp.2.weight", "mlp.4.weight"]
    assert "_tf32_weights" not in mlp.state_dict()


def test_host_memory_near_gpu_restores_the_default_policy():
    """parallel.host_memory_near_gpu: without NUMA information (no GPU here) the block changes nothing; with a node it sets
    MPOL_PREFERRED for the block and puts MPOL_DEFAULT back afterwards (checked with get_mempolicy)."""
    import ctypes
    import platform
    from rq_vae_recommender_b200 import parallel
    with parallel.host_memory_near_gpu() as h:
        assert h.node is None
    nr_get = {"x86_64": 239, "aarch64": 236}.get(platform.machine())
    if nr_get is None:
        return
    libc = ctypes.CDLL(None, use_errno=True)

    def policy():
        mode = ctypes.c_int(-1)
        assert libc.syscall(nr_get, ctypes.byref(mode), None, 0, None, 0) == 0
        return mode.value

    base = policy()
    h = parallel.host_memory_near_gpu()
    h.node = 0                                   # node 0 exists on every Linux host
    with h:
        inside = policy() if h.node is not None else None
    assert policy() == base == 0                 # MPOL_DEFAULT
    assert inside in (1, None)                   # MPOL_PREFERRED, or the syscall is not permitted in this sandbox
