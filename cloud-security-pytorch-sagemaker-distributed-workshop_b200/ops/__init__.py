"""Native ops: C-ABI library binding (`_lib`), allocation-explicit wrappers (`native`), autograd
functions with CPU fallback (`functional`) and fused optimizers (`optim`)."""
