"""Feed / stop rules of the reference's seq2seq Helpers (modules/helpers.py; tacotron2 helpers, SURVEY.md
A.7) as the (feed_mode, stop_mode) pair the persistent decoder kernel takes.  The per-step Helper
callbacks of TF (initialize / sample / next_inputs) are folded into the kernel; these classes keep the
reference's names and constructor arguments and expose which kernel mode they select."""
from .. import _lib


class _Helper:
    feed_mode = None
    stop_mode = None

    def __init__(self, targets, batch_size, output_dim, r, n_feed_frame=1):
        assert n_feed_frame <= r
        self.targets, self.batch_size, self.output_dim, self.r, self.n_feed_frame = \
            targets, batch_size, output_dim, r, n_feed_frame

    @property
    def num_steps(self):
        return None if self.targets is None else self.targets.shape[1] // self.r


class TransformerTrainingHelper(_Helper):
    """modules/helpers.py:13-58: teacher-forced feed, fixed S steps, argmax sample ids."""
    feed_mode, stop_mode = _lib.FEED_TEACHER, _lib.STOP_FIXED

    def __init__(self, targets, output_dim, r, n_feed_frame=1):
        super().__init__(targets, targets.shape[0], output_dim, r, n_feed_frame)


class OneHotValidationHelper(_Helper):
    """modules/helpers.py:61-108: teacher forcing or softmax-probability feedback; finished at S steps."""
    stop_mode = _lib.STOP_FIXED

    def __init__(self, targets, batch_size, output_dim, r, n_feed_frame=1, teacher_forcing=False):
        super().__init__(targets, batch_size, output_dim, r, n_feed_frame)
        self.teacher_forcing = teacher_forcing
        self.feed_mode = _lib.FEED_TEACHER if teacher_forcing else _lib.FEED_SELF_SOFTMAX


class ValidationHelper(_Helper):
    """tacotron2 ValidationHelper (A.7): like OneHotValidationHelper without the softmax."""
    stop_mode = _lib.STOP_FIXED

    def __init__(self, targets, batch_size, output_dim, r, n_feed_frame=1, teacher_forcing=False):
        super().__init__(targets, batch_size, output_dim, r, n_feed_frame)
        self.teacher_forcing = teacher_forcing
        self.feed_mode = _lib.FEED_TEACHER if teacher_forcing else _lib.FEED_SELF


class StopTokenBasedInferenceHelper(_Helper):
    """tacotron2 StopTokenBasedInferenceHelper (A.7; in-tree twin modules/helpers.py:111-160):
    feed back the last n_feed_frame outputs; finished = reduce_all(sigmoid(stop) > 0.5 and time > min_iters)."""
    feed_mode, stop_mode = _lib.FEED_SELF, _lib.STOP_ALL_ROWS

    def __init__(self, batch_size, output_dim, r, n_feed_frame=1, min_iters=10, dtype=None):
        super().__init__(None, batch_size, output_dim, r, n_feed_frame)
        self.min_iters = min_iters
