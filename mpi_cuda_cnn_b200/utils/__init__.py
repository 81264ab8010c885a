"""IDX files, synthetic data, device timing and clock sampling."""
from .idx import read_idx, write_idx, make_synthetic, write_synthetic_dataset  # noqa: F401
from .timing import ClockSampler, cuda_time_ms, flush_l2  # noqa: F401
