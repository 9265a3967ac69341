"""B200-native cut-FEM assembly hot path (ngsxfem drop-in for that path)."""
