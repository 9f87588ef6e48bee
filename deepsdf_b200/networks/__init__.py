"""Mirror of the reference package ``deep_sdf.networks``: ``NetworkArch`` names resolve here."""
