"""FusedTPColumnwise / FusedTPRowwise with tile=wide on one GPU (world 1), validated."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch  # noqa: E402

from ddlb_b200.primitives.registry import load_impl_class  # noqa: E402

for prim, opts in (("tp_columnwise", dict(order="AG_before", tile="wide")),
                   ("tp_columnwise", dict(order="AG_after", tile="wide")),
                   ("tp_rowwise", dict(tile="wide"))):
    impl = load_impl_class(prim, "fused")(m=1024, n=512, k=512, dtype="bfloat16", **opts)
    out = impl.run()
    torch.cuda.synchronize()
    impl.validate(out)
    print(prim, opts, "wide =", impl.op.plan.info["wide"], "ok", flush=True)
