"""ActNorm (reference: normflow/flows/normalization.py:8-38)."""
import torch

from .affine.coupling import AffineConstFlow


class ActNorm(AffineConstFlow):
    """AffineConstFlow whose s, t are set from the first batch it sees, in either direction
    (normalization.py:20-38: unbiased std + 1e-6).  The one-off statistics use two library
    reductions; every later call is the `vcnf_const_affine` kernel."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.data_dep_init_done_cpu = torch.tensor(0.)
        self.register_buffer("data_dep_init_done", self.data_dep_init_done_cpu)

    def initialised(self):
        """True once the data-dependent initialisation has run (or was loaded).  A flag that lives on the device
        (module moved / state_dict loaded after the initialisation) is read once and then kept on the host like the
        reference keeps it after initialising (normalization.py:27,37): no device synchronisation per call."""
        flag = self.data_dep_init_done
        if flag.is_cuda:
            if not bool(flag > 0.):
                return False
            self.data_dep_init_done = torch.tensor(1.)
            return True
        return bool(flag > 0.)

    def forward(self, z):
        if not self.initialised():
            assert self.s is not None and self.t is not None
            with torch.no_grad():
                s_init = -torch.log(z.std(dim=self.batch_dims, keepdim=True) + 1e-6)
                self.s.data = s_init.data.to(self.s.dtype)
                self.t.data = (-z.mean(dim=self.batch_dims, keepdim=True) * torch.exp(self.s)).data.to(self.t.dtype)
            self.data_dep_init_done = torch.tensor(1.)
        return super().forward(z)

    def inverse(self, z):
        if not self.initialised():
            assert self.s is not None and self.t is not None
            with torch.no_grad():
                self.s.data = torch.log(z.std(dim=self.batch_dims, keepdim=True) + 1e-6).data.to(self.s.dtype)
                self.t.data = z.mean(dim=self.batch_dims, keepdim=True).data.to(self.t.dtype)
            self.data_dep_init_done = torch.tensor(1.)
        return super().inverse(z)
