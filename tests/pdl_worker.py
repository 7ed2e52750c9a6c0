"""Worker for test_gpu_tc.test_pdl_on_off_agree: one bf16 decoder fwd + bwd + reverse at the BASELINE shape, results to an
.npz.  GTK_PDL is read once per process by the library, hence a subprocess per setting."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from glowtts_b200 import commons, models  # noqa: E402


def main(out_path):
    cfg = dict(helpers.BASE_CFG, n_blocks=4)
    sd = helpers.make_state_dict(cfg, 91, end_std=0.01)
    dec = models.FlowSpecDecoder(cfg["in_channels"], cfg["hidden_channels"], cfg["kernel_size"], cfg["dilation_rate"],
                                 cfg["n_blocks"], cfg["n_layers"], n_split=cfg["n_split"], n_sqz=cfg["n_sqz"], precision="bf16")
    dec.load_state_dict(sd, strict=True)
    dec = dec.cuda().train()
    y, mask, _, _ = helpers.flow_inputs(cfg, 32, 860, 13, False)
    y, mask = y.cuda().requires_grad_(True), mask.cuda()
    out = {}
    for rep in range(3):                       # repeated: a race would not hit the same way every time
        dec.zero_grad(set_to_none=True)
        y.grad = None
        z, ld = dec(y, mask)
        loss = commons.mle_loss(z, None, None, ld, mask)
        loss.backward()
        with torch.no_grad():
            yb, _ = dec(z.detach(), mask, reverse=True)
        torch.cuda.synchronize()
        out[f"z{rep}"] = z.detach().cpu().numpy()
        out[f"ld{rep}"] = ld.detach().cpu().numpy()
        out[f"dy{rep}"] = y.grad.cpu().numpy()
        out[f"yb{rep}"] = yb.cpu().numpy()
        g = torch.cat([p.grad.reshape(-1) for p in dec.parameters() if p.grad is not None])
        out[f"g{rep}"] = g.cpu().numpy()
    np.savez(out_path, **out)


if __name__ == "__main__":
    main(sys.argv[1])
