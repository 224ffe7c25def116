"""conceptattention_b200 — B200-native ConceptAttention hot path.

Drop-in for the reference's `concept_attention` package surface
(concept_attention/__init__.py:1): `ConceptAttentionFluxPipeline` with
`generate_image` / `encode_image` returning `ConceptAttentionPipelineOutput`.
All device maths on the path goes through the C-ABI library
`libconceptattention_b200.so` (include/conceptattention_b200.h); there is no CPU
or eager fallback — ops raise if the library is missing.
"""
__all__ = ["ConceptAttentionFluxPipeline", "ConceptAttentionPipelineOutput",
           "compute_heatmaps_from_vectors"]


def __getattr__(name):          # lazy: importing .weights must not need the .so
    if name in __all__:
        from . import pipeline
        return getattr(pipeline, name)
    raise AttributeError(name)
