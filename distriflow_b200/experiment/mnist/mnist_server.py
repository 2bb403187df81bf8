"""Server main of the MNIST experiment.

Parity: `/root/reference/experiment/mnist/mnist_server.ts:1-37` — dense model `Flatten -> Dense(10, relu) ->
Dense(10, softmax)`, in-memory server model, asynchronous SGD server with a server-owned dataset, an upload hook that
logs the worker's metrics, listening on a port (the reference uses 80).  The reference's hook prints `metrics[0]` as both
loss and accuracy; here loss and accuracy are printed from their own slots.

    python -m distriflow_b200.experiment.mnist.mnist_server --port 8080 [--mode async|sync] [--workers K]
"""
from __future__ import annotations

import argparse
import time

from ... import AsynchronousSGDServer, DistriServerInMemoryModel, SynchronousSGDServer, TcpServerEndpoint
from ...models import mnist_mlp
from .mnist_data import load_dataset


def create_dense_model():
    return mnist_mlp()


def main(argv=None) -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--port", type=int, default=8080)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--mode", default="async", choices=["async", "sync"])
    ap.add_argument("--workers", type=int, default=2, help="sync mode: uploads per version")
    ap.add_argument("--loss", default="referenceCompat")
    ap.add_argument("--max-staleness", type=int, default=4)
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--synthetic-examples", type=int, default=60000)
    a = ap.parse_args(argv)
    dataset = load_dataset(config={"batchSize": 32, "epochs": a.epochs, "smallLastBatch": False},
                           n_synthetic=a.synthetic_examples)
    model = DistriServerInMemoryModel(create_dense_model(), {"loss": a.loss})
    endpoint = TcpServerEndpoint(a.host, a.port)
    if a.mode == "async":
        server = AsynchronousSGDServer(endpoint, model, dataset, {"verbose": True,
                                                                 "serverHyperparams": {"maximumStaleness": a.max_staleness}})
    else:
        server = SynchronousSGDServer(endpoint, model, dataset, {"verbose": True,
                                                                "serverHyperparams": {"minUpdatesPerVersion": a.workers}})
    server.on_upload(lambda msg: msg and msg.metrics and print(f"loss: {msg.metrics[0]:.5f}, accuracy: {msg.metrics[-1]:.4f}",
                                                                 flush=True))
    server.setup()
    print(f"listening on {endpoint.url}", flush=True)
    try:
        while not server.done:
            time.sleep(0.2)
        time.sleep(0.5)
    finally:
        endpoint.close()
    print(f"done: {server.numUpdates} updates, {server.numDropped} dropped, {len(server.workerIDs)} workers", flush=True)


if __name__ == "__main__":
    main()
