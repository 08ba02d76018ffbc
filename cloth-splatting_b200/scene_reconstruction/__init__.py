"""Hot-path part of the reference's `scene_reconstruction` package: the mesh -> Gaussian transform."""
