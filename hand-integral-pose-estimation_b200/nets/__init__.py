"""Drop-in for the reference's `common/nets` package as far as the integral-regression hot
path goes: `from nets import loss` / `from nets.loss import softmax_integral_tensor, ...`."""
