"""The example app (reference main.py) and its data plumbing, on CPU."""
import os
import subprocess
import sys

import torch

from pipe_b200.utils.data import Vocab, basic_english, batchify, build_corpus, get_batch, synthetic_corpus

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tokenizer_and_vocab(tmp_path):
    assert basic_english("Hello, World!  (twice)") == ["hello", ",", "world", "!", "(", "twice", ")"]
    p = tmp_path / "c.txt"
    p.write_text("the cat sat .\n\nthe dog sat .\n")
    ids, vocab = build_corpus(str(p))
    assert vocab.itos[0] == "<unk>" and len(vocab) == 6
    assert ids.tolist() == vocab("the cat sat . the dog sat .".split())
    assert vocab(["zebra"]) == [0]


def test_batchify_and_get_batch():
    data = torch.arange(26)
    b = batchify(data, 4)
    assert b.shape == (6, 4) and b[:, 0].tolist() == [0, 1, 2, 3, 4, 5] and b[0].tolist() == [0, 6, 12, 18]
    x, y = get_batch(b, 0, 3)
    assert x.shape == (4, 3)                      # batch-first for the pipe
    assert x[0].tolist() == [0, 1, 2] and y[:4].tolist() == [1, 7, 13, 19]
    x, y = get_batch(b, 4, 3)                     # tail: shorter sequence
    assert x.shape == (4, 1)


def test_synthetic_corpus_is_zipfian():
    s = synthetic_corpus(20000, 1000, seed=1)
    assert s.min() >= 0 and s.max() < 1000
    counts = torch.bincount(s, minlength=1000)
    assert counts[0] > counts[10] > counts[200]


def test_example_app_runs_on_cpu():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "main.py"), "except_last", "--small", "--steps", "3"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "Total parameters in model" in out.stdout and "| batch   2 |" in out.stdout


def test_multi_process_example_two_pipelines_of_two_stages_on_cpu():
    """examples/dist_main.py under torchrun with gloo: 2 pipelines x 2 stages.  Also the regression test for
    process-group creation: every rank must create every pipeline's boundary groups (a pipeline that only
    created its own deadlocked the other one on gloo)."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "4", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "examples", "dist_main.py"), "except_last", "--stages", "2",
           "--small", "--steps", "3"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=420)
    assert out.returncode == 0, (out.stdout + out.stderr)[-3000:]
    assert "2 pipeline(s) x 2 stages" in out.stdout
