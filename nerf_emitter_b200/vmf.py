"""vMF-mixture emitter sampling / pdf on the GPU (DS/emitters/vMF.py:24-81), emitter-local coordinates."""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as L


class VmfMixture:
    """Lobe table of ``VonMisesFisherEmitter``: position (K,3), weight (K), std (K)
    (loaded by ``load_vmf_params``, NS/path_guiding/vmf_guiding.py:52-57)."""

    def __init__(self, position: torch.Tensor, weight: torch.Tensor, std: torch.Tensor):
        self.lib = L.load()
        if not position.is_cuda:
            raise L.NemitError("VmfMixture needs CUDA tensors; there is no CPU path")
        K = position.shape[0]
        self.position = position.detach().float().contiguous()
        self.weight = weight.detach().float().contiguous().to(position.device)
        self.std = std.detach().float().contiguous().to(position.device)
        if self.position.shape != (K, 3) or self.weight.shape != (K,) or self.std.shape != (K,):
            raise ValueError("lobe table shapes must be (K,3), (K), (K)")
        self._lobes = L.NemitVmfLobes(self.position.data_ptr(), self.weight.data_ptr(), self.std.data_ptr(), K)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.position.device).cuda_stream)

    def sample_direction(self, p: torch.Tensor, u: torch.Tensor):
        """p (N,3) shading points, u (N,2) in [0,1)^2 -> d (N,3) unit directions, pdf (N)."""
        n = p.shape[0]
        p = p.float().contiguous()
        u = u.float().contiguous()
        d = torch.empty(n, 3, device=p.device, dtype=torch.float32)
        pdf = torch.empty(n, device=p.device, dtype=torch.float32)
        L.check(self.lib.nemit_vmf_sample(C.byref(self._lobes), p.data_ptr(), u.data_ptr(), n, d.data_ptr(),
                                          pdf.data_ptr(), self._stream()), "nemit_vmf_sample")
        return d, pdf

    def pdf_direction(self, p: torch.Tensor, d: torch.Tensor):
        n = p.shape[0]
        p = p.float().contiguous()
        d = d.float().contiguous()
        pdf = torch.empty(n, device=p.device, dtype=torch.float32)
        L.check(self.lib.nemit_vmf_pdf(C.byref(self._lobes), p.data_ptr(), d.data_ptr(), n, pdf.data_ptr(),
                                       self._stream()), "nemit_vmf_pdf")
        return pdf
