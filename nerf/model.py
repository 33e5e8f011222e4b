"""`from nerf.model import NeRF` -- the B200-native module behind the reference's import path.

One `NeRF` name serves the three API generations of the reference: constructor
keywords of the checkpoint generation (`normalize_position`, `density_inputs`, ...) build the G1/G2
model of examples/nerf.pth; the keywords of the class in the reference tree (`color_outputs`,
`segmentation_outputs`, `encoding_size`, `focal_length`, `min_x` ... `max_z`) -- or no arguments at
all, the reference's own default construction -- build the HEAD (mip-NeRF style) model."""
from nerf_b200.head import HeadNeRF
from nerf_b200.model import NeRF as G1NeRF

_HEAD_KEYS = {"color_outputs", "segmentation_outputs", "encoding_size", "focal_length",
              "min_x", "min_y", "min_z", "max_x", "max_y", "max_z"}


class _NeRFMeta(type(G1NeRF)):
    def __call__(cls, *args, **kwargs):
        if cls is NeRF:
            head = (not args and not kwargs) or bool(_HEAD_KEYS & set(kwargs))
            return HeadNeRF(*args, **kwargs) if head else G1NeRF(*args, **kwargs)
        return super().__call__(*args, **kwargs)

    def __instancecheck__(cls, obj):
        return isinstance(obj, (G1NeRF, HeadNeRF))


class NeRF(G1NeRF, metaclass=_NeRFMeta):
    """Factory type: `NeRF(...)` returns the G1/G2 or the HEAD module (see module docstring); the static
    helpers shared by all generations are reachable on the class itself."""
