// encode_kernels.cu -- sm_100a kernels of the b200jpeg encode path (filled in with the encoder milestone).
namespace b200jpeg {
}
