class TorchTrainer:
    def __init__(self, *a, **k):
        raise RuntimeError("ray stub: TorchTrainer is not available")


class TorchConfig:
    def __init__(self, *a, **k):
        self.kwargs = k
