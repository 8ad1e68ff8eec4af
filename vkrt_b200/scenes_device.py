"""Device-side generator of BASELINE.json config 5: the 50 M-triangle procedural soup is far too
large to build with numpy and push through PCIe every frame, so its vertex buffer is produced by torch ops on the
GPU (plumbing) and handed to vkrt_blas_build / vkrt_blas_rebuild with VKRT_BUILD_INPUT_ON_DEVICE.

Layout and distribution are those of scenes.procedural_soup: one geometry, Model::Vertex (64 B) records,
centroids uniform in [-20,20]x[0,15]x[-12,12], edge vectors ~N(0, 0.02^2) per axis, frame f moves centroid.y by
0.05*sin(0.1 f + phase(tri)).  The random stream is torch's (seed 0x5EED), not numpy's: the small-size parity
tests use the numpy generator, this one feeds the full-size benchmark."""
from typing import Dict

import numpy as np

from . import scenes


class SoupDevice:
    def __init__(self, triangles: int, device: int = 0, seed: int = 0x5EED, chunk: int = 1 << 22):
        import torch

        self.torch = torch
        self.n = int(triangles)
        self.dev = torch.device("cuda", device)
        g = torch.Generator(device=self.dev)
        g.manual_seed(seed)
        n = self.n
        self.vertices = torch.zeros((3 * n, 16), dtype=torch.float32, device=self.dev)
        self.base_y = torch.empty((n,), dtype=torch.float32, device=self.dev)
        self.phase = torch.empty((n,), dtype=torch.float32, device=self.dev)
        self.edge_y = torch.empty((n, 2), dtype=torch.float32, device=self.dev)
        scale = torch.tensor([40.0, 15.0, 24.0], device=self.dev)
        shift = torch.tensor([-20.0, 0.0, -12.0], device=self.dev)
        v = self.vertices.view(n, 3, 16)
        for a in range(0, n, chunk):  # chunked so the temporaries stay small next to the 9.6 GB vertex buffer
            b = min(n, a + chunk)
            m = b - a
            c = torch.rand((m, 3), generator=g, device=self.dev) * scale + shift
            e = torch.randn((m, 2, 3), generator=g, device=self.dev) * 0.02
            self.phase[a:b] = torch.rand((m,), generator=g, device=self.dev) * (2.0 * np.pi)
            self.base_y[a:b] = c[:, 1]
            self.edge_y[a:b] = e[:, :, 1]
            nrm = torch.cross(e[:, 0], e[:, 1], dim=1)
            nrm = nrm / torch.clamp(torch.linalg.norm(nrm, dim=1, keepdim=True), min=1e-20)
            tng = e[:, 0] / torch.clamp(torch.linalg.norm(e[:, 0], dim=1, keepdim=True), min=1e-20)
            v[a:b, 0, 0:3] = c
            v[a:b, 1, 0:3] = c + e[:, 0]
            v[a:b, 2, 0:3] = c + e[:, 1]
            v[a:b, :, 4:7] = nrm[:, None, :]
            v[a:b, :, 12:15] = tng[:, None, :]
            v[a:b, :, 15] = 1.0
        self.indices = torch.arange(3 * n, dtype=torch.int32, device=self.dev)
        self.set_frame(0)

    def set_frame(self, frame: int):
        """Moves every triangle to its position in animation frame `frame` (in place, positions only)."""
        torch = self.torch
        y = self.base_y + 0.05 * torch.sin(0.1 * float(frame) + self.phase)
        v = self.vertices.view(self.n, 3, 16)
        v[:, 0, 1] = y
        v[:, 1, 1] = y + self.edge_y[:, 0]
        v[:, 2, 1] = y + self.edge_y[:, 1]
        torch.cuda.synchronize(self.dev)  # the library copies on its own stream
        return self.vertices

    def ranges(self) -> np.ndarray:
        return np.array([[self.n, 3 * self.n - 1, 0]], dtype=np.uint64)

    def model_stub(self) -> Dict:
        """Materials/images of config 5 (white 4x4 texture, material row {0,0,0,0}); geometry stays on the device."""
        return {"materials": [(0, 0, 0)], "images": [scenes.solid_image((255, 255, 255, 255), 4)],
                "provenance": f"procedural soup, {self.n} triangles, generated on the device (torch, seed 0x5EED)"}
