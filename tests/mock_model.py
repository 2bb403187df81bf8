"""MockModel: isolates the protocol from any real math (reference: src/test/mock_model.ts:5-46)."""
import torch

from distriflow_b200 import DistriModel
from distriflow_b200.checkpoint import new_version


class MockModel(DistriModel):
    isDistributedClientModel = True
    isDistributedServerModel = True
    inputShape = [1]
    outputShape = [1]

    def __init__(self, vars_):
        super().__init__()
        self.vars = [v.clone() for v in vars_]
        self.updates = 0

    def setup(self):
        pass

    def save(self):
        self.version = new_version()
        self.versionIndex += 1

    def fit(self, x, y):
        return self.vars          # "gradients" = own variables

    def update(self, grads):
        self.updates += 1
        self.last_grads = [g.clone() for g in grads]

    def set_vars(self, vals):
        for v, n in zip(self.vars, vals):
            v.copy_(n)

    def get_vars(self):
        return self.vars

    def predict(self, x):
        return x

    def evaluate(self, x, y):
        return [0.0]
