// Engine: backward pass (filled in incrementally).
#include "engine.h"

namespace act {

int Engine::backward(int stage, float loss_grad, cudaStream_t st) {
  (void)stage; (void)loss_grad; (void)st;
  set_error("backward not implemented yet");
  return -1;
}

}  // namespace act
