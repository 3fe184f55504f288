from ecolab_model import panmictic_ecolab as ecolab
ecolab.density([1, 0, 1, 0, 1])
ecolab.makeConsistent()
ecolab.repro_rate([1, 1, 1, 1, 1])
ecolab.interaction.diag([1, 1, 1, 1, 1])
ecolab.condense()
assert ecolab.density()==[1, 1, 1]
