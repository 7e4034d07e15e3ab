"""CPU oracle of the hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; it is the checker, never the product.  The product
(imaginaire_b200/) never imports it and fails loudly when its CUDA library is
missing.

spade_oracle.py   SPADE layers, residual block, generator (fp32 torch-CPU restatement; QUANT / STORE emulation hooks)
train_oracle.py   SPADE discriminator, loss terms, dis_update / gen_update / moving average of one training iteration
warp_oracle.py, resample_oracle.c          the two flow warps (numpy driver + plain C)
flownet_ops_oracle.{py,c}                  FlowNet2 channelnorm / correlation
ref_cuda/                                  the reference's own CUDA kernel bodies compiled for the host (-> oracle/_ref/)
make_golden*.py                            fixture generators that import the reference in the build container
"""
