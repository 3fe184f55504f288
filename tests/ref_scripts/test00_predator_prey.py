from ecolab_model import panmictic_ecolab as ecolab
from ecolab import array_urand

ecolab.seed(10)
ecolab.species([1, 2])
ecolab.density([100, 100]) 
ecolab.create([0, 0])
ecolab.repro_rate([.1, -.1])
ecolab.interaction.diag([-.0001, -1e-5])
ecolab.interaction.val([-0.001, 0.001])
ecolab.interaction.row([0, 1])
ecolab.interaction.col([1, 0])
ecolab.mutation([.01, .01])
ecolab.sp_sep(.1)
ecolab.repro_min(-.1)
ecolab.repro_max(.1)
ecolab.odiag_min(-1e-3)
ecolab.odiag_max(1e-3)
ecolab.mut_max(.01)

with open("out.dat","w") as output:
     for j in range(10):
        for i in range(100):
            ecolab.generate()
        ecolab.condense()
        d=ecolab.density()
        print(d[0],d[1], file=output)

