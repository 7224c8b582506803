"""Where does the mel_post error come from?  CPU-only attribution with the oracle PostNet (model/model.py:114-135): bf16
rounding of the PostNet input / activations / weights one at a time, and the gain the PostNet applies to an input error.
Usage: python tools/postnet_error_attribution.py  (output in profiles/r03_postnet_error_attribution.txt)"""
import sys, torch
import os; sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import tts_oracle as O
torch.set_num_threads(8)
sd=O.make_state_dict(41)
ph,mel,pl,ml=O.make_inputs(8,150,800,seed=42)
ref=O.forward(sd,ph,mel,pl,ml)
pre=ref["pred_melspec"]
r=O.bf16_round
def post(x, act_round=False, w_round=False):
    sd2=dict(sd)
    if w_round:
        for k in sd:
            if k.startswith("postnet") and k.endswith("conv.weight"): sd2[k]=r(sd[k])
    for i in range(5):
        x=O.conv_bn(sd2,f"postnet.layers.{3*i}",x)
        if i<4:
            x=torch.tanh(x)
            if act_round: x=r(x)
    return x
exact=post(pre)
for name,x,a,w in (("bf16 input only",r(pre),False,False),("bf16 activations only",pre,True,False),("bf16 weights only",pre,False,True),
                   ("hi+lo input, bf16 act+w",pre,True,True),("bf16 input+act+w",r(pre),True,True)):
    y=post(x,a,w)
    print(f"{name:28s} max_abs={float((y-exact).abs().max()):.4f} rms={float((y-exact).pow(2).mean().sqrt()):.5f}")
g=torch.Generator().manual_seed(0)
for s in (0.005,0.01,0.02):
    n=torch.randn(pre.shape,generator=g)*s
    y=post(pre+n)
    print(f"input + N(0,{s})              out rms={float((y-exact).pow(2).mean().sqrt()):.5f} (x{float((y-exact).pow(2).mean().sqrt())/s:.2f}) max_abs={float((y-exact).abs().max()):.4f} (input max_abs {float(n.abs().max()):.4f})")
