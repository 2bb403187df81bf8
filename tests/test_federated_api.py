"""The reference's three integration scenarios (src/test/federated_api_test.ts:56-90) on the new API,
first over the in-process hub, then over real TCP sockets."""
import threading

import pytest
import torch

from distriflow_b200 import FederatedClient, FederatedServer, LoopbackHub, TcpServerEndpoint
from mock_model import MockModel

INIT = [torch.tensor([[1.0, 1], [1, 1]]), torch.tensor([[1.0, 2, 3, 4]])]


def make(tmp_path, endpoint, client_server):
    server_model = MockModel(INIT)
    server_model.version = "initial"
    client_model = MockModel([torch.zeros_like(t) for t in INIT])
    server = FederatedServer(endpoint, server_model, {"modelDir": str(tmp_path),
                                                      "serverHyperparams": {"minUpdatesPerVersion": 2},
                                                      "clientHyperparams": {"examplesPerUpdate": 1}})
    server.setup()
    client = FederatedClient(client_server, client_model)
    client.setup()
    return server, client, client_model


@pytest.fixture(params=["loopback", "tcp"])
def rig(request, tmp_path):
    if request.param == "loopback":
        hub = LoopbackHub()
        server, client, cm = make(tmp_path, hub, hub.client)
        yield server, client, cm
        client.dispose()
    else:
        ep = TcpServerEndpoint()
        server, client, cm = make(tmp_path, ep, ep.url)
        yield server, client, cm
        client.dispose()
        ep.close()


def test_transmits_model_version_on_startup(rig):
    server, client, cm = rig
    assert client.model_version() == "initial"
    assert torch.equal(cm.vars[1], INIT[1])           # weights arrived too


def test_transmits_updates(rig):
    server, client, cm = rig
    assert len(server.updates) == 0
    cm.vars[0].copy_(torch.full((2, 2), 2.0))
    client.distributed_update(torch.tensor([[0.0]]), torch.tensor([[0.0]]))
    assert len(server.updates) == 1
    assert client.clientId in server.workerIDs


def test_triggers_a_download_after_enough_uploads(rig):
    server, client, cm = rig
    seen = []
    done = threading.Event()

    def cb(old, new):
        seen.append((old, new))
        done.set()

    client.on_new_version(cb)
    cm.vars[0].copy_(torch.full((2, 2), 2.0))
    client.distributed_update(torch.tensor([[0.0]]), torch.tensor([[0.0]]))
    cm.vars[0].copy_(torch.ones(2, 2))
    cm.vars[1].copy_(torch.tensor([[5.0, 4, 3, 1]]))
    client.distributed_update(torch.zeros(3, 1), torch.zeros(3, 1))
    if hasattr(client.socket, "hub"):
        client.socket.hub.pump()
    assert done.wait(5.0)
    old, new = seen[0]
    assert old == "initial" and new != "initial" and new == server.model.version
    # mean of the two buffered uploads went into update()
    assert torch.allclose(server.model.last_grads[0], torch.full((2, 2), 1.5))
    # uploads for the stale version are dropped (federated_server.ts:73)
    assert server.numDropped >= 1 or client.num_updates() >= 2


def test_hyperparam_precedence(rig):
    server, client, cm = rig
    assert client.num_examples_per_update() == 1          # server-sent beats default 5
    client.hyperparams["examplesPerUpdate"] = 3           # local beats server-sent
    assert client.num_examples_per_update() == 3
    client.distributed_update(torch.zeros(2, 1), torch.zeros(2, 1))
    assert client.num_examples() == 2 and client.num_examples_remaining() == 1
