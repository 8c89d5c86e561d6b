"""spectavi_b200 -- B200-native build of spectavi's descriptor-matching path.

``spectavi_b200.feature`` mirrors the NN section of the reference's
``spectavi/feature.py`` (same function names, arguments and return shapes);
``spectavi_b200.cndarray`` is the Python-3 NdArray marshalling;
``spectavi_b200.device`` is the device-resident API over torch tensors used
for sharded (one process per GPU) matching and by bench.py.
"""
__version__ = "0.1.0"
