from .meta_container import MetaContainer, build_expert  # noqa: F401
from .meta_ngp import MetaNGP  # noqa: F401
from .meta_vanilla import MetaNeRF  # noqa: F401,E402
