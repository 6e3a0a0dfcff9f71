"""TEST INFRASTRUCTURE ONLY: CPU restatements of the reference's string path
(py_oracle.py, kx_oracle.c), the loader that executes the reference's own app.py
(ref_loader.py) and the fp32 CPU model used to check the generator.  The product
package never imports anything from here."""
