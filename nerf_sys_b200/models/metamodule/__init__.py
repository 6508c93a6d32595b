from .metamodule import MetaLayerBlock, MetaLinear, MetaModule, MetaSequential  # noqa: F401
