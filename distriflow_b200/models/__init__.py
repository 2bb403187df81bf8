"""Model zoo: the reference's MNIST models (`/root/reference/experiment/mnist/mnist_server.ts:16-22`,
`/root/reference/experiment/mnist/model.json`) plus the BASELINE.json workloads (ResNet-20/50, GPT-2 small)."""
from .topology import SequentialModel, model_from_topology, sequential, topology_from_model
from .mnist import mnist_cnn, mnist_mlp
