"""Rigid frames container with the reference's ``T`` surface used on the path
(genie/utils/affine_utils.py:54-126,243-244): ``T(rots, trans)``, ``.rots``, ``.trans``,
``.shape``, indexing, ``scale_translation``.  Frame math itself runs inside the CUDA
kernels; this class only carries the two tensors across the Python boundary.
"""
import torch


class T:
    def __init__(self, rots, trans):
        if rots is None and trans is None:
            raise ValueError('Only one of rots and trans can be None')
        if rots is None:
            rots = torch.eye(3, dtype=trans.dtype, device=trans.device).expand(*trans.shape[:-1], 3, 3).contiguous()
        if trans is None:
            trans = torch.zeros(*rots.shape[:-2], 3, dtype=rots.dtype, device=rots.device)
        if rots.shape[-2:] != (3, 3) or trans.shape[-1] != 3 or rots.shape[:-2] != trans.shape[:-1]:
            raise ValueError('Incorrectly shaped input')
        self.rots = rots
        self.trans = trans

    def __getitem__(self, index):
        if type(index) != tuple:
            index = (index,)
        return T(self.rots[index + (slice(None), slice(None))], self.trans[index + (slice(None),)])

    @property
    def shape(self):
        s = self.rots.shape[:-2]
        return s if len(s) > 0 else torch.Size([1])

    def get_trans(self):
        return self.trans

    def get_rots(self):
        return self.rots

    def scale_translation(self, factor):
        return T(self.rots, self.trans * factor)

    def to(self, device):
        return T(self.rots.to(device), self.trans.to(device))
