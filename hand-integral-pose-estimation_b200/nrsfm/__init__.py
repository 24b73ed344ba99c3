"""Drop-in for the SVD-backed pieces of the reference's procrustes_encoding/nrsfm package."""
