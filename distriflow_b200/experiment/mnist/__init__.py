"""The MNIST experiment of the reference (`/root/reference/experiment/mnist/`): data loader, server main, client main,
plus the B200 multi-GPU run of the same experiment."""
