"""Python faces of the sm_100a kernels: autograd functions (functional), the tensor-core conv
dispatcher (tc_conv), the multi-tensor optimizer launch (sgd) and the graph-captured ConvNet
training engine (convnet_engine)."""
