"""ugu_b200 — B200-native (sm_100a) drop-in for ONE hot path of unclearness/ugu: the software mesh
rasterizer `ugu::Rasterizer` behind the `ugu::RendererCpu` interface.

Layout: csrc/ (CUDA kernels + the C-ABI of include/ugu_b200_raster.h), cpp/ (the C++
`ugu::RasterizerCuda : ugu::RendererCpu` adapter compiled against the reference headers),
renderer.py (the same interface for Python hosts), multiview.py (view sharding over ranks),
meshgen.py (synthetic inputs).  There is no CPU fallback: importing the renderer without the built
CUDA library raises.
"""
from .renderer import (ColorInterpolation, DiffuseColor, DiffuseShading, GBuffer, Image1b, Image1f, Image1i,
                       Image1w, Image3b, Image3f, Material, Mesh, PinholeCamera, Rasterizer, Raytracer,
                       RendererCpuOption, ShadingNormal)

__all__ = ["ColorInterpolation", "DiffuseColor", "DiffuseShading", "GBuffer", "Image1b", "Image1f", "Image1i",
           "Image1w", "Image3b", "Image3f", "Material", "Mesh", "PinholeCamera", "Rasterizer", "Raytracer",
           "RendererCpuOption", "ShadingNormal"]
