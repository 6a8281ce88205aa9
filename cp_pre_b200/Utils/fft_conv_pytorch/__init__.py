from cp_pre_b200.Utils.fft_conv_pytorch.fft_conv import FFTConv1d, FFTConv2d, FFTConv3d, fft_conv  # noqa: F401
