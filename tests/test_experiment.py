"""The MNIST experiment mains (reference: experiment/mnist/*) over real TCP, and the IDX loader."""
import os
import struct
import threading

import numpy as np
import pytest
import torch

from distriflow_b200.experiment.mnist import mnist_client, mnist_data, mnist_server


def _write_idx(dirpath, n=64):
    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, size=(n, 28, 28), dtype=np.uint8)
    labels = rng.integers(0, 10, size=(n,), dtype=np.uint8)
    for name_i, name_l in (mnist_data.TRAIN, mnist_data.TEST):
        with open(os.path.join(dirpath, name_i), "wb") as f:
            f.write(struct.pack(">IIII", 0x803, n, 28, 28) + imgs.tobytes())
        with open(os.path.join(dirpath, name_l), "wb") as f:
            f.write(struct.pack(">II", 0x801, n) + labels.tobytes())
    return imgs, labels


def test_idx_loader(tmp_path):
    imgs, labels = _write_idx(str(tmp_path))
    m = mnist_data.load_mnist(str(tmp_path))
    x, y = m["train"]
    assert x.shape == (64, 28, 28, 1) and x.dtype == torch.float32 and float(x.max()) > 1.0      # raw 0..255 (Q4)
    assert np.array_equal(x[:, :, :, 0].numpy().astype(np.uint8), imgs) and np.array_equal(y.numpy(), labels)
    ds = mnist_data.load_dataset(str(tmp_path))
    assert ds.batchSize == 32 and ds.epochs == 5 and ds.batches == 2 and ds.y.shape == (64, 10)
    open(os.path.join(tmp_path, mnist_data.TRAIN[0]), "r+b").write(struct.pack(">I", 0x1234))
    with pytest.raises(ValueError, match="invalid magic"):
        mnist_data.load_mnist(str(tmp_path))


def test_server_and_two_clients_over_tcp():
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    st = threading.Thread(target=mnist_server.main, args=(["--port", str(port), "--epochs", "1", "--synthetic-examples", "256"],))
    st.start()
    import time
    time.sleep(0.5)
    cs = [threading.Thread(target=mnist_client.main, args=(["--server", f"tcp://127.0.0.1:{port}", "--client-id", f"w{i}"],))
          for i in range(2)]
    for c in cs:
        c.start()
    for c in cs:
        c.join(60)
    st.join(60)
    assert not st.is_alive() and not any(c.is_alive() for c in cs)


def test_bench_reference_arm_reports_unavailable():
    """`bench.py --impl reference` contract: one JSON line with `"impl": "reference"` and `"unavailable"`, exit code 0
    (the reference is TypeScript on tf.js / socket.io and cannot be installed offline)."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and isinstance(line["unavailable"], str) and line["unavailable"]
