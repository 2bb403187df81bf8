"""distriflow_b200 — a Blackwell-native data-parallel training framework with DistriFlow's capabilities.

Public surface (names of the reference in parentheses):
  DistriServer (AsynchronousSGDServer / FederatedServer), DistriWorker (AsynchronousSGDClient /
  FederatedClient), DistriModel (DistributedModel family), DistriDataset (DistributedDataset),
  checkpoint IO in the tf.js LayersModel layout, and — new for B200 — `parallel.DataParallelEngine`,
  where every worker is a GPU rank and the server is sharded across the ranks.
"""
from .config import (DEFAULT_CLIENT_HYPERPARAMS, DEFAULT_DATASET_HYPERPARAMS, DEFAULT_DISTRIBUTED_COMPILE_ARGS,
                     DEFAULT_ENGINE_CONFIG, DEFAULT_SERVER_HYPERPARAMS, ConfigError, client_hyperparams,
                     compile_args, dataset_config, engine_config, override, server_hyperparams)
from .serialization import (SerializedVariable, deserialize_var, deserialize_vars, flat_deserialize, flat_serialize,
                            serialize_var, serialize_vars, serialized_to_array, stack_serialized)
from .messages import Batch, DataMsg, DownloadMsg, Events, GradientMsg, ModelMsg, UploadMsg
from .dataset import DistriDataset, DistributedDataset, batch_to_data_msg
from .model import (DistriDynamicModel, DistriModel, DistriTorchModel, DistributedDynamicModel, DistributedModel,
                    DistributedTfModel, StalenessPolicy)
from .checkpoint import (DistriServerDynamicModel, DistriServerInMemoryModel, DistriServerTorchModel,
                         DistributedServerDynamicModel, DistributedServerInMemoryModel, DistributedServerTfModel,
                         fetch_model, is_distributed_server_model, load_layers_model, save_layers_model)
from .transport import LoopbackHub, TcpClientEndpoint, TcpServerEndpoint
from .server import AbstractServer, AsynchronousSGDServer, DistriServer, FederatedServer, SynchronousSGDServer
from .worker import (AbstractClient, AsynchronousSGDClient, DistriClientTorchModel, DistriWorker, FederatedClient,
                     is_distributed_client_model)
from .trainer import Trainer
from .b200 import B200Model, B200Server, B200Worker

__version__ = "0.1.0"
