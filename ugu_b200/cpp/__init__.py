"""C++ host side of the drop-in: ugu::RasterizerCuda : ugu::RendererCpu (rasterizer_cuda.{h,cc}),
compiled against the reference headers, plus its differential harness (adapter_check.cc)."""
