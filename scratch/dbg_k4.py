import sys, torch
sys.path.insert(0, '.')
import lt_b200
from oracle import ref_ops
for shape in [(64,17,96,96),(2,2,6,8),(2,3,224,224),(1,2,256,256),(3,5,13,31)]:
    gen = torch.Generator().manual_seed(7)
    hm0 = torch.randn(*shape, generator=gen); mult = 100.0
    h64 = (hm0.double()*mult).requires_grad_(True)
    c_ref, p_ref = ref_ops.integrate_tensor_2d(h64, True)
    gc = torch.randn(c_ref.shape, generator=gen)
    (g_ref,) = torch.autograd.grad(c_ref, h64, gc.double(), retain_graph=True)
    hm = hm0.cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_2d(hm, True, multiplier=mult)
    (gh,) = torch.autograd.grad(coords, hm, gc.cuda(), retain_graph=True)
    ref = g_ref*mult; d = (gh.cpu().double()-ref).abs()
    flat = d.view(-1).argmax().item()
    idx = torch.unravel_index(torch.tensor(flat), d.shape)
    b, j = idx[0].item(), idx[1].item()
    print(shape, 'rel', (d.max()/ref.abs().max()).item(), 'at', [i.item() for i in idx], 'ours', gh.cpu()[idx].item(), 'ref', ref[idx].item(), 'maxref', ref.abs().max().item())
    # per-map relative error
    dm = d.flatten(2).max(2).values; rm = ref.abs().flatten(2).max(2).values
    print('   worst per-map rel', (dm/rm).max().item(), ' map with max |ref|:', rm.view(-1).argmax().item(), ' map with max err', dm.view(-1).argmax().item())
    pm = p_ref[b, j].flatten().topk(3)
    print('   top3 p of that map', pm.values.tolist(), 'ours', prob[b, j].flatten().topk(3).values.tolist())
