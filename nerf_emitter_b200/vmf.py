"""vMF-mixture emitter sampling / pdf on the GPU (DS/emitters/vMF.py:24-81), emitter-local coordinates."""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as L


class VmfMixture:
    """Lobe table of ``VonMisesFisherEmitter``: position (K,3), weight (K), std (K)
    (loaded by ``load_vmf_params``, NS/path_guiding/vmf_guiding.py:52-57)."""

    def __init__(self, position: torch.Tensor, weight: torch.Tensor, std: torch.Tensor, world_transform=None,
                 bsphere_center=(0.5, 0.5, 0.5), bsphere_radius: float = 0.8660254037844386):
        """world_transform: the emitter's to_world as a (4,4) / (3,4) matrix (None = identity: points and directions are
        emitter-local); bsphere: the emitter's bounding sphere m_bsphere (default: the unit cube's)."""
        self.lib = L.load()
        self.set_world_transform(world_transform, bsphere_center, bsphere_radius)
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

    def set_world_transform(self, world_transform=None, bsphere_center=(0.5, 0.5, 0.5), bsphere_radius: float = 0.8660254037844386):
        """params['to_world'] of the emitter (Rotater.apply_sdf_scene moves it per rotation, rotater.py:115-123)."""
        T = torch.eye(4, dtype=torch.float64)
        if world_transform is not None:
            W = torch.as_tensor(world_transform, dtype=torch.float64).cpu()
            T[:W.shape[0], :4] = W[:, :4]
        Ti = torch.linalg.inv(T)
        f = L.NemitVmfFrame()
        for i, x in enumerate(Ti[:3, :4].reshape(-1).tolist()):
            f.world_to_local[i] = x
        for i, x in enumerate(T[:3, :4].reshape(-1).tolist()):
            f.local_to_world[i] = x
        for i in range(3):
            f.bsphere_center[i] = float(bsphere_center[i])
        f.bsphere_radius = float(bsphere_radius)
        self._frame = f
        self.world_transform = T
        return self

    def sample_direction_world(self, it_p: torch.Tensor, u: torch.Tensor):
        """VonMisesFisherEmitter.sample_direction (vMF.py:24-56) for world-space reference points it_p (N,3):
        -> DirectionSample fields {d (N,3), pdf (N), dist (N), p (N,3), n (N,3) = -d}; the returned weight is 0."""
        n = it_p.shape[0]
        it_p = it_p.float().contiguous()
        u = u.float().contiguous()
        f = dict(device=it_p.device, dtype=torch.float32)
        d, pdf, dist, p = torch.empty(n, 3, **f), torch.empty(n, **f), torch.empty(n, **f), torch.empty(n, 3, **f)
        L.check(self.lib.nemit_vmf_sample_direction(C.byref(self._lobes), C.byref(self._frame), it_p.data_ptr(), u.data_ptr(), n,
                                                    d.data_ptr(), pdf.data_ptr(), dist.data_ptr(), p.data_ptr(), self._stream()),
                "nemit_vmf_sample_direction")
        return {"d": d, "pdf": pdf, "dist": dist, "p": p, "n": -d}

    def pdf_direction_world(self, it_p: torch.Tensor, ds_d: torch.Tensor):
        """VonMisesFisherEmitter.pdf_direction (vMF.py:64-81) for world-space it.p and ds.d."""
        n = it_p.shape[0]
        it_p = it_p.float().contiguous()
        ds_d = ds_d.float().contiguous()
        pdf = torch.empty(n, device=it_p.device, dtype=torch.float32)
        L.check(self.lib.nemit_vmf_pdf_direction(C.byref(self._lobes), C.byref(self._frame), it_p.data_ptr(), ds_d.data_ptr(), n,
                                                 pdf.data_ptr(), self._stream()), "nemit_vmf_pdf_direction")
        return pdf

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


N_CLUSTER = 64     # NS/path_guiding/vmf_guiding.py:18


class VmfGuiding:
    """Host mirror of ``VonMisesFisherGuiding`` (NS/path_guiding/vmf_guiding.py:60-118) and of the ``PathGuiding``
    hooks (NS/path_guiding/path_guiding.py:20-75) for the lobe tables this path consumes: holds the primal emitter's
    mixture (and any adjoint emitters'), (de)serialises them under the reference's ``state_dict`` keys
    ``vmf.{position,weight,std}`` / ``adjoint_emitter_i.*`` and exposes the no-op training hooks the op calls."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.mixtures = {}

    def load_vmf_params(self, name: str, cluster):
        """load_vmf_params (vmf_guiding.py:52-57): cluster = {position (K,3), weight (K), std (K)}."""
        self.mixtures[name] = VmfMixture(torch.as_tensor(cluster["position"]).to(self.device).reshape(-1, 3),
                                         torch.as_tensor(cluster["weight"]).to(self.device).reshape(-1),
                                         torch.as_tensor(cluster["std"]).to(self.device).reshape(-1))
        return self.mixtures[name]

    @property
    def primal(self) -> VmfMixture:
        return self.mixtures["vmf"]

    def state_dict(self, destination=None, prefix=""):
        destination = {} if destination is None else destination
        for name, m in self.mixtures.items():
            destination[prefix + name + ".position"] = m.position.reshape(-1)      # dr.ravel order: xyzxyz...
            destination[prefix + name + ".weight"] = m.weight
            destination[prefix + name + ".std"] = m.std
        return destination

    def load_pipeline(self, loaded_state, step: int = 0):
        names = sorted({k.rsplit(".", 1)[0] for k in loaded_state if k.endswith(".position")
                        and (k.startswith("vmf") or k.startswith("adjoint_emitter_"))})
        for name in names:
            self.load_vmf_params(name, {key: loaded_state.pop(name + "." + key) for key in ("position", "weight", "std")})

    # PathGuiding's hooks called from the op (nerf_emitter_op.py:57,88,117): no-ops for the vMF guiding
    def train_primal(self, *a):
        pass

    def train_forward(self, *a):
        pass

    def train_backward(self, *a):
        pass
