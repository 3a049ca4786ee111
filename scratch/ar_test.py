import os, torch, torch.distributed as dist, time
rank=int(os.environ['RANK']); world=int(os.environ['WORLD_SIZE']); lr=int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(lr)
dist.init_process_group('nccl', device_id=torch.device('cuda',lr))
P=500000
buf=torch.randn(P,10,device='cuda')
def timeit(fn, n=20):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    s=torch.cuda.Event(enable_timing=True); e=torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e)/n*1e3
t=timeit(lambda: dist.all_reduce(buf))
if rank==0: print(f'nccl all_reduce 20MB: {t:.1f} us')
try:
    import torch.distributed._symmetric_memory as symm
    t2 = symm.empty(P*10, dtype=torch.float32, device='cuda')
    hdl = symm.rendezvous(t2, dist.group.WORLD)
    t2.normal_()
    for name in ['one_shot_all_reduce','two_shot_all_reduce_','multimem_all_reduce_']:
        try:
            op = getattr(torch.ops.symm_mem, name)
            tt=timeit(lambda: op(t2, 'sum', dist.group.WORLD.group_name))
            if rank==0: print(f'symm_mem {name}: {tt:.1f} us')
        except Exception as ex:
            if rank==0: print(name, 'failed', repr(ex)[:200])
    if rank==0: print('has multicast', hdl.multicast_ptr if hasattr(hdl,'multicast_ptr') else None, 'buffer_ptrs', [hex(p) for p in hdl.buffer_ptrs][:2], 'signal', len(hdl.signal_pad_ptrs))
except Exception as ex:
    if rank==0: print('symm_mem unavailable', repr(ex)[:300])
dist.destroy_process_group()
