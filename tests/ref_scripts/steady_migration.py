from ecolab_model import spatial_ecolab as ecolab

from random import random, seed as randomSeed

from ecolab import array_urand, myid, device

print(device())

# we want initialisation to be identical across all processes
randomSeed(10)

# we want the array operations to have a different seed across processes
array_urand.seed(10+myid())

# initial number of species
nsp=100

ecolab.repro_min(-0.1)
ecolab.repro_max(0.1)
ecolab.odiag_min(-1e-4)
ecolab.odiag_max(1e-4)
#ecolab.mut_max(1e-4)
ecolab.mut_max(1e-3)
ecolab.sp_sep(0.1)
ecolab.gen_bias(0.0)
ecolab.fixMigration(True)

def randomList(num, min, max):
    return [random()*(max-min)+min for i in range(num)]

ecolab.species(range(nsp))

numX=2
numY=2
ecolab.setGrid(numX,numY)
ecolab.partitionObjects()

print("initialising density")
for i in range(numX):
    for j in range(numY):
        #ecolab.cell(i,j).density([int(100*random()) for i in range(nsp)])
        ecolab.cell(i,j).density(nsp*[100])
print("density initialised")
        
ecolab.repro_rate(randomList(nsp, ecolab.repro_min(), ecolab.repro_max()))
ecolab.interaction.diag(randomList(nsp, -1e-3, -1e-3))
ecolab.random_interaction(3,0)
              
ecolab.interaction.val(randomList(len(ecolab.interaction.val), ecolab.odiag_min(), ecolab.odiag_max()))

ecolab.mutation(nsp*[ecolab.mut_max()])
ecolab.migration(nsp*[1e-3])
                  
from plot import plot
from GUI import gui, statusBar, windows

epoch=1000000
mut_factor=1000

extinctions=0
migrations=0
savedMig=ecolab.migration()
def stepImpl():
    ecolab.generate(100)
    ecolab.mutate()

    global savedMig

    global extinctions, migrations
    migrations+=ecolab.migrate()
    extinctions+=ecolab.condense()

print(ecolab.nsp()())
ecolab.makeConsistent()
ecolab.syncThreads()
print(ecolab.nsp()())

out=open("steady.dat","w")
print('#tstep','nsp','connectivity','migrations','extinctions',file=out,flush=True)

def step():
    global extinctions,migrations
    extinctions=0
    migrations=0
    for i in range(epoch//10000):
        stepImpl()
    #print('migrations=',migrations,' extinctions=',extinctions)
    if myid()==0:
        nsp=len(ecolab.species)
        print('tstep=',ecolab.tstep(),'nsp=',nsp,'migrations=',migrations,' extinctions=',extinctions)
        print(ecolab.tstep(),nsp,ecolab.connectivity(),migrations,extinctions,file=out,flush=True)
        statusBar.configure(text=f't={ecolab.tstep()} nsp:{nsp}')
#        plot('No. species',ecolab.tstep(),nsp,200*(ecolab.tstep()%epoch<0.5*epoch))
#        #plot('No. species',ecolab.tstep(),nsp)
#        plot('No. species by cell',ecolab.tstep(),ecolab.nsp()())
#        plot('Extinctions',ecolab.tstep(),extinctions)
#        plot('Migration',ecolab.tstep(),migrations)
#        for i in range(numX):
#            for j in range(numY):
#                plot(f'Density({i},{j})',ecolab.tstep(),ecolab.cell(i,j).density(), pens=ecolab.species())

#gui(step)

while ecolab.tstep()<5e6:
    step()



