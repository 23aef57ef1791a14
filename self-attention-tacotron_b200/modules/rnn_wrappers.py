"""Names of /root/reference/modules/rnn_wrappers.py.  In the reference these are TF RNNCell
compositions evaluated once per `while_loop` iteration; here the whole composition
(OutputAndStopTokenTransparentWrapper(TransformerWrapper(RNNStateHistoryWrapper(DecoderRNNV2(...)))),
SURVEY.md 3.2) is ONE persistent CUDA kernel, so the wrappers are declarative: they record the
composition and `compose()` checks that it is the one the kernel implements."""


class TransparentRNNCellLike:
    """modules/rnn_wrappers.py:18-39."""

    def __init__(self, cell=None):
        self._cell = cell

    @property
    def cell(self):
        return self._cell

    def layers(self):
        inner = self._cell.layers() if isinstance(self._cell, TransparentRNNCellLike) else [type(self._cell).__name__]
        return [type(self).__name__] + inner


class RNNStateHistoryWrapper(TransparentRNNCellLike):
    """:47-80 -- appends each decoder output to the history (here: the KV cache rows in HBM)."""

    def __init__(self, cell, max_iter):
        super().__init__(cell)
        self.max_iter = max_iter


class TransformerWrapper(TransparentRNNCellLike):
    """:87-124 -- re-attends over the history each step; the kernel evaluates the same last row from a
    KV cache (identical because of the causal mask)."""

    def __init__(self, cell, transformers, memory_sequence_length):
        super().__init__(cell)
        self.transformers = transformers
        self.memory_sequence_length = memory_sequence_length


class OutputAndStopTokenTransparentWrapper(TransparentRNNCellLike):
    """:188-214 -- frame projection [B, M*r] and stop-token projection [B, 1]."""

    def __init__(self, cell, out_units, out_projection, stop_token_projection):
        super().__init__(cell)
        self.out_units = out_units
        self.out_projection, self.stop_token_projection = out_projection, stop_token_projection


def compose(cell):
    """Wrapper stack -> the persistent-kernel form that implements it: 'DualSourceTransformerDecoder' (KV-cached
    self-attention hops inside the step) or None for the plain projection-wrapped cell (ExtendedDecoder /
    DualSourceDecoder differ in their attention mechanisms, not in the cell stack); anything else raises."""
    names = cell.layers()
    if names[:3] == ["OutputAndStopTokenTransparentWrapper", "TransformerWrapper", "RNNStateHistoryWrapper"]:
        return "DualSourceTransformerDecoder"
    if names == ["OutputAndStopTokenTransparentWrapper", "str"]:
        return None  # plain DecoderRNNV2 under the projection wrapper: ExtendedDecoder / DualSourceDecoder
    raise ValueError(f"unsupported decoder cell composition: {names}")
