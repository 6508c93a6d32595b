from .meta_container import MetaContainer, build_expert  # noqa: F401
from .meta_ngp import MetaNGP  # noqa: F401
