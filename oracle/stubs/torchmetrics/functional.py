"""Import stub (see ../README.md): SSIM is an evaluation metric, off the per-ray path."""


def structural_similarity_index_measure(*a, **k):
    raise RuntimeError("stub torchmetrics.functional.structural_similarity_index_measure")
