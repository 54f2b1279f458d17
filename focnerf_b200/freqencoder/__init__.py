from .freq import FreqEncoder, freq_encode, _backend  # noqa: F401
