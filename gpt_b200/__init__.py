"""gpt_b200 — B200-native hot path of hyunjik11/GPT (tensor-RFF GP regression sampled by SGLD).

`gpt_b200.gpt_sgld` keeps the names and argument order of the reference's Julia module GPT_SGLD
(/root/reference/GPT_SGLD.jl:5) and calls libgptb200.so through its C ABI (include/gptb200.h).
There is no CPU fallback: importing works anywhere, computing needs a B200.
"""
from . import gpt_sgld  # noqa: F401
from .gpt_sgld import (  # noqa: F401
    GPTregression, RFFFeatures, datawhitening, feature, feature2, geod, pred, pred_samples, proj, samplenz,
    GeodesicNaN, GptbError, GPNT_SGLD, GPT_SGLDERMw, NTFeatures, featureNotensor, pred_notensor,
)
