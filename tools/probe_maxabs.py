import sys, torch
sys.path.insert(0,'.')
import lora_phy_b200 as lp
phy=lp.LoraPhyCuda(0); dev=phy.device
for sf,L,P in ((7,32,65536),(12,64,256)):
    cfg=lp.make_cfg(sf)
    g=torch.Generator().manual_seed(1)
    pay=torch.randint(0,256,(P,L),dtype=torch.uint8,generator=g)
    sym=phy.encode_batch(pay.to(dev)); iq=phy.modulate_batch(cfg,sym); phy.dechirp_batch(cfg,iq)
    torch.cuda.synchronize()
    m=torch.view_as_real(iq).abs().amax(dim=(1,2))
    print(sf,"packets with max>1:",int((m>1).sum()),"of",P,"max",float(m.max()))
