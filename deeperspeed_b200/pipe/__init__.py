from ..runtime.pipe import PipelineModule, LayerSpec, TiedLayerSpec
