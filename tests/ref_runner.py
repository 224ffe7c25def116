"""The comparator of the end-to-end parity tests: the reference's OWN modules (oracle/_ref,
built by oracle/make_ref.py) when they are present, else the oracle port — on any device
and in any dtype.  bf16 on CUDA = the reference's real numerics on this GPU (cuBLAS GEMMs +
the library SDPA kernel torch picks on sm_100); fp32 = exact maths (TF32 is off by default
for matmuls in torch)."""
import torch

import oracle
from oracle import reference_modules as rm

PAIRS = (("output_space_image_vectors", "output_space_concept_vectors"),
         ("cross_attention_image_vectors", "cross_attention_concept_vectors"))


class RefRunner:
    def __init__(self, params, state_dict, device, dtype):
        self.p, self.device, self.dtype = params, torch.device(device), dtype
        if rm.available():
            self.kind = "reference modules (" + rm.load().kind + ")"
            self.model = rm.build_dit(params.asdict(), state_dict, self.device, dtype)
        else:                                     # pragma: no cover  (oracle/_ref travels with the snapshot)
            self.kind = "oracle port"
            self.sd = {k: v.to(device=self.device, dtype=dtype) for k, v in state_dict.items()}

    def _cast(self, inp):
        return {k: v.to(device=self.device, dtype=self.dtype) for k, v in inp.items()}

    @torch.no_grad()
    def denoise(self, inp, sched, guidance=0.0, joint_attention_kwargs=None):
        """Euler loop over `sched` -> (latent tokens, {key: [Nt,L,B,...]})
        (flux/src/flux/sampling.py:96-152)."""
        ii = self._cast(inp)
        if hasattr(self, "model"):
            x, _, st = rm.load().denoise(
                self.model, img=ii["img"], img_ids=ii["img_ids"], txt=ii["txt"], txt_ids=ii["txt_ids"],
                vec=ii["vec"], timesteps=sched, guidance=guidance, concepts=ii["concepts"],
                concept_ids=ii["concept_ids"], concept_vec=ii["concept_vec"],
                joint_attention_kwargs=joint_attention_kwargs)
            return x, st
        return oracle.denoise(self.sd, self.p.asdict(), ii["img"], ii["img_ids"], ii["txt"], ii["txt_ids"],
                              ii["vec"], sched, guidance=guidance, concepts=ii["concepts"],
                              concept_ids=ii["concept_ids"], concept_vec=ii["concept_vec"],
                              joint_attention_kwargs=joint_attention_kwargs)

    @torch.no_grad()
    def forward(self, inp, t, guidance=None, **kw):
        """One DiT forward -> (pred | None, {key: [L,B,...]}) (modified_flux_dit.py:75-163)."""
        ii = self._cast(inp)
        t = t.to(device=self.device, dtype=self.dtype)
        g = None if guidance is None else guidance.to(device=self.device, dtype=self.dtype)
        if hasattr(self, "model"):
            return self.model(img=ii["img"], img_ids=ii["img_ids"], txt=ii["txt"], txt_ids=ii["txt_ids"],
                              concepts=ii["concepts"], concept_ids=ii["concept_ids"],
                              concept_vec=ii["concept_vec"], y=ii["vec"], timesteps=t, guidance=g, **kw)
        return oracle.dit_forward(self.sd, self.p.asdict(), ii["img"], ii["img_ids"], ii["txt"], ii["txt_ids"],
                                  ii["concepts"], ii["concept_ids"], ii["concept_vec"], t, ii["vec"], g, **kw)

    @torch.no_grad()
    def heatmaps(self, stacked, layers, timesteps, **kw):
        """(concept_heatmaps, cross_attention_maps) [B,C,h,w] in the run's dtype
        (concept_attention_pipeline.py:29-91,150-169)."""
        if hasattr(self, "model"):
            return rm.heatmaps(stacked, layers, timesteps, **kw)
        return tuple(oracle.heatmaps_from_vectors(stacked[a], stacked[b], list(layers), list(timesteps), **kw)
                     for a, b in PAIRS)


def rel_map(got, ref):
    """max |got - ref| / max(|ref|, 1e-2): north_star's "max relative error" with a floor that
    keeps near-zero probabilities from dominating."""
    got, ref = torch.as_tensor(got).float().cpu(), torch.as_tensor(ref).float().cpu()
    return ((got - ref).abs() / ref.abs().clamp_min(1e-2)).max().item()


def mae(got, ref):
    got, ref = torch.as_tensor(got).float().cpu(), torch.as_tensor(ref).float().cpu()
    return (got - ref).abs().mean().item()


def rel_max(got, ref):
    got, ref = torch.as_tensor(got).float().cpu(), torch.as_tensor(ref).float().cpu()
    return ((got - ref).abs().max() / ref.abs().max()).item()


def mask_agreement(a, b):
    """Fraction of pixels whose argmax over the concept axis (dim -3) agrees."""
    a, b = torch.as_tensor(a).float().cpu(), torch.as_tensor(b).float().cpu()
    return (a.argmax(-3) == b.argmax(-3)).float().mean().item()
