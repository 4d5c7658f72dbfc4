import sys, time, torch, cProfile, pstats
sys.path.insert(0, "tests")
import parity as T
m = T.load_ours()
case = T.make_case(camera="tum", P=20000, view_k=1)
S = T.settings_for(m, case, "cuda")
dev = {k: case[k].cuda() for k in ("means3D","opacities","shs","scales","rotations")}
Wc, Wd = case["Wc"].cuda(), case["Wd"].cuda()
leaves = {k: v.clone().requires_grad_(True) for k, v in dev.items()}
m2 = torch.zeros(20000, 3, device="cuda", requires_grad=True)
theta = torch.zeros(3, device="cuda", requires_grad=True); rho = torch.zeros(3, device="cuda", requires_grad=True)
def run():
    c, r, d, o, n = m.GaussianRasterizer(S)(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"], theta=theta, rho=rho)
    torch.autograd.backward([c, d], [Wc, Wd])
for _ in range(20): run()
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(200): run()
torch.cuda.synchronize(); print("ours 20k: %.3f ms/it" % ((time.perf_counter() - t) * 5))
ref = T.load_ref()
Sr = T.settings_for(ref, case, "cuda")
def runr():
    c, r, d, o, n = ref.GaussianRasterizer(Sr)(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"], theta=theta, rho=rho)
    torch.autograd.backward([c, d], [Wc, Wd])
for _ in range(20): runr()
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(200): runr()
torch.cuda.synchronize(); print("ref 20k: %.3f ms/it" % ((time.perf_counter() - t) * 5))
pr = cProfile.Profile(); pr.enable()
for _ in range(200): run()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
