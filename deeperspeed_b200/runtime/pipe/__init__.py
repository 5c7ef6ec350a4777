from .module import PipelineModule, LayerSpec, TiedLayerSpec
from .topology import ProcessTopology
