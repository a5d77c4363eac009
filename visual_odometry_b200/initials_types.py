"""Row types of the hot path (reference src/initials.py:16-33), same field names and order."""
import collections

Feature = collections.namedtuple('Feature', ['keypoint', 'descriptor', 'feature_id'])
_MATCH_FIELDS = ['featureid1', 'featureid2', 'keypoint1', 'keypoint2',
                 'descriptor1', 'descriptor2', 'distance', 'color']
Match = collections.namedtuple('Match', _MATCH_FIELDS)
Match3D = collections.namedtuple('Match3D', _MATCH_FIELDS + ['point'])
MatchWithMap = collections.namedtuple(
    'MatchWithMap', ['featureid1', 'featureid2', 'imagecoord', 'mapcoord',
                     'descriptor1', 'descriptor2', 'distance'])
