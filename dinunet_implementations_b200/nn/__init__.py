from .trainer import COINNTrainer, NNTrainer

__all__ = ["COINNTrainer", "NNTrainer"]
