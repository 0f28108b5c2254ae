from parameter_server_b200.utils.range import Range, U64_MAX


def test_even_divide_covers_range():
    r = Range(3, 1003)
    parts = [r.even_divide(7, i) for i in range(7)]
    assert parts[0].begin == 3 and parts[-1].end == 1003
    assert all(a.end == b.begin for a, b in zip(parts, parts[1:]))


def test_even_divide_all_uint64():
    parts = [Range.all().even_divide(3, i) for i in range(3)]
    assert parts[0].begin == 0 and parts[-1].end == U64_MAX
    assert all(a.end == b.begin for a, b in zip(parts, parts[1:]))
    assert abs(parts[0].size() - U64_MAX // 3) <= 2


def test_intersection_and_contains():
    a, b = Range(0, 10), Range(5, 20)
    assert a.intersection(b) == Range(5, 10) and a.contains(9) and not a.contains(10)
    assert Range(5, 5).empty()
