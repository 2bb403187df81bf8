"""Client (worker) main of the MNIST experiment.

Parity: `/root/reference/experiment/mnist/mnist_client.ts:1-31` — same dense model, an asynchronous SGD client that
connects to the server, sends pre-step metrics and loops download -> fit -> upload until the server stops sending work.

    python -m distriflow_b200.experiment.mnist.mnist_client --server tcp://127.0.0.1:8080 [--device cuda]
"""
from __future__ import annotations

import argparse

import torch

from ... import AsynchronousSGDClient, DistriClientTorchModel
from ...models import mnist_mlp


def main(argv=None) -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--server", default="tcp://127.0.0.1:8080")
    ap.add_argument("--loss", default="referenceCompat")
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--client-id", default=None)
    a = ap.parse_args(argv)
    model = DistriClientTorchModel(mnist_mlp(), {"loss": a.loss}, device=torch.device(a.device))
    client = AsynchronousSGDClient(a.server, model, {"sendMetrics": True, "clientId": a.client_id})
    client.setup()
    n = 1 + client.run(idle_timeout=3.0)
    print(f"worker {client.clientId}: {n} micro-batches, {client.num_versions()} model versions", flush=True)
    client.dispose()


if __name__ == "__main__":
    main()
