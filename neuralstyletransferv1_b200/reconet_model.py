"""``lib.ReCoNetModel`` (lib.py:15-64) on the B200 path: uint8 NHWC numpy in, uint8 NHWC numpy out."""
from __future__ import annotations

import numpy as np
import torch

from .modules import ReCoNet


class ReCoNetModel:
    def __init__(self, state_dict_path, use_gpu=True, gpu_device=None, frn=False):
        if not use_gpu:
            raise RuntimeError("ReCoNetModel (B200 drop-in) has no CPU path; use_gpu must be True")
        self.device = torch.device("cuda", 0 if gpu_device is None else int(gpu_device))
        self.model = ReCoNet(frn=frn)
        state = state_dict_path if isinstance(state_dict_path, dict) else torch.load(state_dict_path, map_location="cpu")
        self.model.load_state_dict(state)
        self.model = self.model.to(self.device)
        self.model.eval()

    def run(self, images):
        """lib.py:27-51: /255 -> *2-1 -> model -> (y+1)/2 -> clamp(y*255,0,255).to(uint8)."""
        assert images.dtype == np.uint8
        assert 3 <= images.ndim <= 4
        orig_ndim = images.ndim
        if images.ndim == 3:
            images = images[None, ...]
        x = torch.from_numpy(np.ascontiguousarray(images)).to(self.device)
        y = self.model.stylize_u8(x, "tanh")
        y = y.cpu().numpy()
        return y[0] if orig_ndim == 3 else y
