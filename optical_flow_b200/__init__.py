"""optical_flow_b200: B200-native (sm_100a) hot path of xianlopez/optical_flow.

Drop-in Python surface (same names as the reference):
  transformations.bilinear_interpolation, model.{warp_features, create_cost_volume,
  upscale_flow, flow_module, reset18_encoder, build_flow_net}, loss.LossLayer,
  train.train_step, old.image_warp.{image_warp, image_warp_tf}, old.loss.MyLoss,
  resnet.models.{resnet_block_simple, resnet_layer_simple}.
All arithmetic runs in liboflow_b200.so (hand-written CUDA behind a C ABI).
"""
__version__ = "0.1.0"
