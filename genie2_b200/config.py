"""Key-value text configuration with the reference's keys and defaults
(genie/config.py:9-107): lines of ``camelCaseKey value``; booleans only when the value is
literally ``True``/``False``.  Produces the same ``io`` / ``diffusion`` / ``model`` /
``training`` / ``optimization`` dictionaries the reference's callers index.
"""

_INT, _FLOAT, _RAW = int, float, (lambda x: x)

# section -> [(dict key, file key, default, cast)]
_TABLE = {
    'io': [
        ('name', 'name', None, _RAW), ('rootdir', 'rootDirectory', 'runs', _RAW),
        ('datadir', 'dataDirectory', 'data/afdbreps_l-256_plddt_80/pdbs', _RAW),
        ('min_n_res', 'minimumNumResidues', 20, _INT), ('max_n_res', 'maximumNumResidues', 256, _INT),
        ('max_n_chain', 'maximumNumChains', 1, _INT), ('validation_split', 'validationSplit', None, _FLOAT),
        ('motif_prob', 'motifProbability', 0.8, _FLOAT), ('motif_min_pct_res', 'motifMinimumPercentageResidues', 0.05, _FLOAT),
        ('motif_max_pct_res', 'motifMaximumPercentageResidues', 0.5, _FLOAT),
        ('motif_min_n_seg', 'motifMinimumNumberSegments', 1, _INT), ('motif_max_n_seg', 'motifMaximumNumberSegments', 4, _INT),
    ],
    'diffusion': [('n_timestep', 'numTimesteps', 1000, _INT), ('schedule', 'schedule', 'cosine', _RAW)],
    'model': [
        ('c_s', 'singleFeatureDimension', 384, _INT), ('c_p', 'pairFeatureDimension', 128, _INT), ('rescale', 'rescale', 1, _FLOAT),
        ('c_pos_emb', 'positionalEmbeddingDimension', 256, _INT), ('c_chain_emb', 'chainEmbeddingDimension', 64, _INT),
        ('c_timestep_emb', 'timestepEmbeddingDimension', 512, _INT),
        ('relpos_k', 'relativePositionK', 32, _INT), ('template_dist_min', 'templateDistanceMinimum', 2, _FLOAT),
        ('template_dist_step', 'templateDistanceStep', 0.5, _FLOAT), ('template_dist_n_bin', 'templateDistanceNumBins', 37, _INT),
        ('n_pair_transform_layer', 'numPairTransformLayers', 5, _INT),
        ('include_mul_update', 'includeTriangularMultiplicativeUpdate', True, _RAW),
        ('include_tri_att', 'includeTriangularAttention', False, _RAW),
        ('c_hidden_mul', 'triangularMultiplicativeHiddenDimension', 128, _INT),
        ('c_hidden_tri_att', 'triangularAttentionHiddenDimension', 32, _INT), ('n_head_tri', 'triangularAttentionNumHeads', 4, _INT),
        ('tri_dropout', 'triangularDropout', 0.25, _FLOAT), ('pair_transition_n', 'pairTransitionN', 4, _INT),
        ('n_structure_layer', 'numStructureLayers', 8, _INT), ('n_structure_block', 'numStructureBlocks', 1, _INT),
        ('c_hidden_ipa', 'ipaHiddenDimension', 16, _INT), ('n_head_ipa', 'ipaNumHeads', 12, _INT),
        ('n_qk_point', 'ipaNumQkPoints', 4, _INT), ('n_v_point', 'ipaNumVPoints', 8, _INT), ('ipa_dropout', 'ipaDropout', 0.1, _FLOAT),
        ('n_structure_transition_layer', 'numStructureTransitionLayers', 1, _INT),
        ('structure_transition_dropout', 'structureTransitionDropout', 0.1, _FLOAT),
    ],
    'training': [
        ('seed', 'seed', 100, _INT), ('n_epoch', 'numEpoches', 1, _INT), ('batch_size', 'batchSize', 1, _INT),
        ('log_every_n_step', 'logEverySteps', 1000, _INT), ('checkpoint_every_n_epoch', 'checkpointEveryEpoches', 500, _INT),
        ('condition_loss_weight', 'conditionLossWeight', 1, _INT),
    ],
    'optimization': [('lr', 'learningRate', 1e-4, _FLOAT)],
}


class Config:

    def __init__(self, filename=None):
        raw = {} if filename is None else self._load_config(filename)
        for section, rows in _TABLE.items():
            out = {}
            for key, fkey, default, cast in rows:
                v = raw.get(fkey, default)
                out[key] = cast(v) if v is not None else None
            setattr(self, section, out)

    @staticmethod
    def _load_config(filename):
        cfg = {}
        with open(filename) as fh:
            for line in fh:
                elts = line.split()
                if len(elts) == 2:
                    cfg[elts[0]] = True if elts[1] == 'True' else False if elts[1] == 'False' else elts[1]
        return cfg
