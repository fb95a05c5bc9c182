"""``tf.keras`` facade: exactly the symbols the reference uses."""
from . import callbacks, datasets, layers, losses, metrics, mixed_precision, models, optimizers  # noqa: F401
from .models import History, Model, Sequential  # noqa: F401


def load_model(filepath, strategy=None):
    from .. import export
    return export.load_model(filepath)


models.load_model = load_model
