"""Postnet (TacotronModel/modules/Postnet.py:5-36): 5 x conv1d k5 (tanh x4, identity) over
[B, F, 80] decoder outputs.  On the B200 path the conv stack, the 768 -> 80 projection, the residual
add and the output clip (Tacotron.py:287-309) are one C-ABI call (``taco_postnet``: tcgen05
implicit-GEMM kernels, bf16 operands / fp32 accumulate, or fp32 SIMT)."""
from .Encoder import FusedStageError


class Postnet(object):
    def __init__(self, is_training, hparams, activation='tanh', scope=None):
        if is_training:
            raise FusedStageError("training-mode postnet (dropout + batch statistics) runs inside train/step.py (Tacotron.initialize(is_training=True))")
        self.is_training = is_training
        self.kernel_size = hparams.postnet_kernel_size
        self.channels = hparams.postnet_channels + hparams.style_embed_depth   # Postnet.py:22
        self.activation = activation
        self.scope = 'postnet_convolutions' if scope is None else scope
        self.postnet_num_layers = hparams.postnet_num_layers
        self.drop_rate = hparams.dropout_rate
        self.hparams = hparams

    def __call__(self, inputs):
        """``inputs``: raw decoder frames [B, F, 80] (device tensor).  Returns
        (decoder_output [B,F,80] clipped, mel_outputs [B,F,80]) -- the residual / projection / clip
        that follow the conv stack in Tacotron.initialize are part of the same fused stage."""
        from ... import runtime
        eng = runtime.engine(self.hparams)
        return eng.postnet(inputs.contiguous(), inputs.shape[1])
