a = 1 + 2 * 3 ^ 2 ^ 0.5 - -4 / 8 % 3
b = (1 + 2) * 3
c = "n=" .. 10 / 4 .. "!"
d = not nil == true
e = {10, 20, 30, [4] = 40, k = {z = -1.5e-3}}
f = #e + e[2] + e.k.z
g = 0x10 + .5 + 3.
h = math.max(1, 7, 3) + math.floor(-0.5) + math.sqrt(16)
i = false or "dflt"
j = nil and 1
k = 2 < 3 and 3 >= 3 and 1 ~= 2
t = {} ; t.u = 5 ; t.v = { w = t.u * 2 }
