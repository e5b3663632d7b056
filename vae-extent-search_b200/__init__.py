"""vae-extent-search_b200: B200-native hot path of Jiw00n/vae-extent-search.

The directory name carries a hyphen, so import it through the `vaex_b200` alias package at the
repo root (`import vaex_b200`), which maps onto this directory."""
from . import _lib  # noqa: F401
from .dataset import Dataset, LearningTask  # noqa: F401
from .mlp_model import (MLPModel, MLPModelInternal, RowsImage, SegmentDataLoader, SegmentSumMLPModule,  # noqa: F401
                        flatten_features, make_net, predict_flat)
from .vae_model import VAE_feature_head, VAECostPredictor, snap_extents, topk  # noqa: F401
from . import extent, feature, select, shard, seg_models, seg_training, training  # noqa: F401
from .seg_models import SegmentVAE  # noqa: F401
from .select import select_programs  # noqa: F401
from .extent import ExtentScaler, SketchSlots, state_to_records  # noqa: F401
